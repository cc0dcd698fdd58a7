#!/usr/bin/env python
"""Every kernel of a slab step, on ONE GPU: `--world` slab engines side by side on device 0
(libclsph_b200.slab.LoopbackWorld: the real pack / unpack / sort ... / commit kernels, messages
moved by a device copy), `--particles` in all, `--steps` steps.  Meant to run under
`ncu --metrics gpu__time_duration.sum` for a launch list of the multi-GPU step's kernels."""
import argparse
import os
import sys
import tempfile

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=8388608)
    ap.add_argument("--world", type=int, default=2)
    ap.add_argument("--steps", type=int, default=3)
    a = ap.parse_args()
    from libclsph_b200 import host
    from libclsph_b200.slab import CudaSlab, LoopbackWorld, equal_count_cuts
    with tempfile.TemporaryDirectory() as tmp:
        case = bench.build_case(tmp, a.particles, "water_box")
    n = case["n"]
    p = host.lattice(n, case["volume"])
    ids = np.arange(n, dtype=np.uint32)
    x = p["position"][:, 0]
    cuts = equal_count_cuts(x, a.world)
    per = int(np.ceil(np.cbrt(n)))
    exch = 3 * per * per * 4 + 65536
    locs = [CudaSlab(case["params"], case["terms"], case["scene"], 0, int(1.2 * n / a.world) + 2 * exch, exch)
            for _ in range(a.world)]
    lw = LoopbackWorld(locs, cuts)
    parts = []
    for r in range(a.world):
        m = (x >= np.float32(cuts[r])) & (x < np.float32(cuts[r + 1]))
        parts.append((p[m], ids[m]))
    lw.start(parts)
    dt = 1.0 / 600.0
    for _ in range(a.steps):
        dt = lw.step(dt)
    print("done", dt, [d.n_owned for d in lw.ranks], [d.n_ghost for d in lw.ranks])


if __name__ == "__main__":
    main()
