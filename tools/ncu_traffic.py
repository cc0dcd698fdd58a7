#!/usr/bin/env python
"""profiles/ncu_traffic.json from an `ncu --set full` report of the two neighbour passes.

    python tools/ncu_traffic.py <prof.ncu-rep> <particles>

Per kernel (density / forces) and particle count: DRAM bytes per launch (dram__bytes_read.sum +
dram__bytes_write.sum), warp and thread instructions executed, L1 data-pipe wavefronts, the launch
duration under ncu — tagged with the hash of the kernel sources they were measured on
(bench.kernel_source_hash) and the commit.  bench.py copies them into `roofline` only while that
hash still matches the sources it runs (a stale capture is reported as null + the reason)."""
import csv
import io
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402

WANT = dict(dram_read="dram__bytes_read.sum", dram_write="dram__bytes_write.sum",
            warp_inst="smsp__inst_executed.sum", thread_inst="smsp__thread_inst_executed.sum",
            lane_efficiency="smsp__thread_inst_executed_per_inst_executed.ratio",
            l1_wavefronts_per_sm="l1tex__data_pipe_lsu_wavefronts.avg", time="gpu__time_duration.sum",
            issue_active_pct="smsp__issue_active.avg.pct_of_peak_sustained_active",
            l1_wavefronts_pct="l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
            warps_active_pct="sm__warps_active.avg.pct_of_peak_sustained_active",
            registers="launch__registers_per_thread")
SCALE = {"Gbyte": 1e9, "Mbyte": 1e6, "Kbyte": 1e3, "byte": 1.0, "ms": 1e-3, "us": 1e-6, "ns": 1e-9, "s": 1.0}


def main(path, n):
    out = subprocess.run(["ncu", "-i", path, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rows = list(csv.reader(io.StringIO(out)))
    head, units = rows[0], rows[1]
    dest = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    doc = dict(source_hash=bench.kernel_source_hash(),
               commit=subprocess.run(["git", "rev-parse", "--short", "HEAD"], capture_output=True, text=True,
                                     cwd=ROOT).stdout.strip(),
               how="ncu --set full --clock-control none, one launch each (tools/round_gpu.sh)", density={}, forces={})
    doc["kernels"] = {}  # every captured kernel (first launch of each): DRAM bytes and duration under ncu
    for r in rows[2:]:
        d = dict(zip(head, r))
        name = d.get("Kernel Name", "")
        short = name.split("(")[0].split("::")[-1].replace("unnamed>", "").strip().split("<")[0]
        if short and short not in doc["kernels"]:
            vals = {}
            for k in ("dram_read", "dram_write", "time"):
                hit = [h for h in head if h == WANT[k] or h.endswith("." + WANT[k])]
                if hit and d[hit[0]] != "":
                    vals[k] = float(d[hit[0]].replace(",", "")) * SCALE.get(units[head.index(hit[0])], 1.0)
            if "time" in vals:
                b = vals.get("dram_read", 0.0) + vals.get("dram_write", 0.0)
                doc["kernels"][short] = dict(particles=n, dram_bytes=b, seconds=vals["time"], dram_gbs=b / vals["time"] / 1e9)
        which = "density" if "k_density" in name else ("forces" if "k_forces_union" in name else None)
        if not which or str(n) in doc[which]:
            continue
        rec = dict(kernel=name.split("(")[0].split("::")[-1].replace("unnamed>", "").strip())
        for k, m in WANT.items():
            hit = [h for h in head if h == m or h.endswith("." + m)]
            if hit and d[hit[0]] != "":
                rec[k] = float(d[hit[0]].replace(",", "")) * SCALE.get(units[head.index(hit[0])], 1.0)
        rec["dram_bytes"] = rec.get("dram_read", 0.0) + rec.get("dram_write", 0.0)
        # (--set full has the per-SM average of the wavefront counter and the lane efficiency, not their sums)
        rec["l1_wavefronts"] = rec.get("l1_wavefronts_per_sm", 0.0) * 148
        rec["thread_inst"] = rec.get("thread_inst") or rec.get("warp_inst", 0.0) * rec.get("lane_efficiency", 32.0)
        doc[which][str(n)] = rec
    json.dump(doc, open(dest, "w"), indent=1)
    print(json.dumps(doc, indent=1))


if __name__ == "__main__":
    main(sys.argv[1], int(sys.argv[2]))
