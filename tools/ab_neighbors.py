#!/usr/bin/env python
"""A/B of the neighbour-pass variants on ONE box: per-pass CUDA-event times of `steps` steps.

    python tools/ab_neighbors.py [--particles N] [--variants 0,2] [--steps K] [--warmup W]
                                 [--kind water_box|dam_break|mucus_monkeybox] [--jitter J] [--evolve S]

variant 0 = the pair-list kernels (k_density_union / k_forces_union), 1 = forces by its own
search, 2 = the round-1 kernels (k_density_pair / k_forces_nlist).  --jitter J displaces every
lattice particle by uniform +-J spacings (mt19937(12345)); --evolve S runs S
steps first, so that the timed state is a flowing fluid and not the rest lattice.  Prints one JSON
line per variant; both variants must agree on the accelerations (max relative difference printed).
"""
import argparse
import json
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=16777216)
    ap.add_argument("--variants", default="0,2")
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--kind", default="water_box")
    ap.add_argument("--jitter", type=float, default=0.0)
    ap.add_argument("--evolve", type=int, default=0)
    ap.add_argument("--options", default="", help="name=value,... applied to every engine")
    a = ap.parse_args()
    from libclsph_b200 import Engine, PASS_NAMES, host
    with tempfile.TemporaryDirectory() as tmp:
        case = bench.build_case(tmp, a.particles, a.kind)
    n, sc = case["n"], case["scene"]
    stage, stage_ptr = bench.pinned_particles(n)
    host.lattice(n, case["volume"], out=stage)
    if a.jitter > 0:
        spacing = float(np.cbrt(case["volume"] / n))
        rng = np.random.Generator(np.random.MT19937(12345))
        stage["position"][:, :3] += ((rng.random((n, 3), dtype=np.float32) * 2 - 1) * np.float32(a.jitter * spacing))
    ref_acc = None
    for v in [int(x) for x in a.variants.split(",")]:
        e = Engine(case["params"], case["terms"], 0)
        e.set_option("neighbor_kernel", v)
        for kv in filter(None, a.options.split(",")):
            k, val = kv.split("=")
            e.set_option(k, int(val))
        e.set_scene(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
        e.upload_ptr(stage_ptr.value, n)
        dt = 1.0 / 600.0
        for _ in range(a.evolve + a.warmup):
            dt = e.step(dt)
        e.set_option("side_stream", 0)
        e.profile(True)
        e.event_record(0)
        for _ in range(a.steps):
            dt = e.step(dt)
        e.event_record(1)
        ms = e.event_elapsed_ms(0, 1) / a.steps
        pass_ms, _, psteps = e.pass_times()
        out = e.download()
        acc = out["acceleration"][:: max(1, n // 200000), :3].astype(np.float64)
        diff = None
        if ref_acc is None:
            ref_acc = acc
        else:
            diff = float(np.max(np.abs(acc - ref_acc)) / max(1e-30, np.max(np.abs(ref_acc))))
        line = dict(variant=v, n=n, kind=a.kind, jitter=a.jitter, evolve=a.evolve, ms_per_step=ms,
                    passes={k: v / max(1, psteps) for k, v in pass_ms.items() if v > 0},
                    acc_rel_diff_vs_first=diff, finite=bool(np.isfinite(acc).all()), dt=dt,
                    lists=e.neighbor_list_stats() if v == 0 else None, list_cap=os.environ.get("SPHB_LIST_CAP"))
        print(json.dumps(line), flush=True)
        e.close()


if __name__ == "__main__":
    main()
