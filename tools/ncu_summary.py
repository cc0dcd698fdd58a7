#!/usr/bin/env python
"""Summarises ncu outputs brought back in gpurun_out/ into small text files for profiles/.

    python tools/ncu_summary.py launches <launches.csv>      per-kernel totals and shares
    python tools/ncu_summary.py kernel <prof.ncu-rep>        key metrics of each captured launch
"""
import collections
import csv
import io
import subprocess
import sys

KEYS = ["gpu__time_duration.sum", "dram__bytes_read.sum", "dram__bytes_write.sum",
        "gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed",
        "sm__throughput.avg.pct_of_peak_sustained_elapsed",
        "l1tex__throughput.avg.pct_of_peak_sustained_elapsed",
        "lts__throughput.avg.pct_of_peak_sustained_elapsed",
        "smsp__issue_active.avg.pct_of_peak_sustained_active",
        "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread",
        "launch__grid_size", "launch__block_size", "smsp__inst_executed.sum",
        "l1tex__t_sector_hit_rate.pct", "lts__t_sector_hit_rate.pct",
        "smsp__average_warp_latency_issue_stalled_long_scoreboard_per_warp_active.pct",
        "smsp__average_warps_issue_stalled_long_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_short_scoreboard_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_lg_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_mio_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_math_pipe_throttle_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_wait_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_barrier_per_issue_active.ratio",
        "smsp__average_warps_issue_stalled_not_selected_per_issue_active.ratio"]


def short(name):
    return name.split("::")[-1].split("(")[0]


def launches(path):
    lines = [l for l in open(path) if l.startswith('"')]
    rows = list(csv.DictReader(io.StringIO("".join(lines))))
    tot, cnt = collections.defaultdict(float), collections.Counter()
    for r in rows:
        k = short(r["Kernel Name"])
        tot[k] += float(r["Metric Value"].replace(",", "")) / 1e6
        cnt[k] += 1
    s = sum(tot.values())
    print(f"# {path}: {len(rows)} launches, {s:.3f} ms total (ncu: cold-cache, serialised; compare shares)")
    for k, v in sorted(tot.items(), key=lambda x: -x[1]):
        print(f"{k:28s} launches {cnt[k]:4d}  total {v:10.3f} ms  avg {1e3 * v / cnt[k]:10.1f} us  share {100 * v / s:5.1f}%")


def kernel(path):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units = rows[0], rows[1]
    for r in rows[2:]:
        d = dict(zip(head, r))
        print(f"## {short(d.get('Kernel Name', '?'))}  grid {d.get('Grid Size')} block {d.get('Block Size')}")
        for k in KEYS:
            if k in d:
                print(f"  {k:90s} {d[k]:>16s} {units[head.index(k)]}")


if __name__ == "__main__":
    {"launches": launches, "kernel": kernel}[sys.argv[1]](sys.argv[2])
