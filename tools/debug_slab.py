"""debug: slab world-1 vs whole-domain engine, per-step diagnostics"""
import sys, os, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from common import make_case, new_engine, state_arrays
from test_gpu_slab import slab_world, reference_run, by_id

world = int(sys.argv[1]) if len(sys.argv) > 1 else 1
with tempfile.TemporaryDirectory() as tmp:
    case = make_case(tmp, "water", 32768)
    ref = reference_run(case, 4)
    lw = slab_world(case, world)
    dt = 1.0 / 600.0
    for s in range(4):
        R = lw.ranks
        dt_in = dt
        dt = lw.step(dt)
        for d in R:
            r = d.local.red8_to_host(d.red8)
            print(f"step {s} dt_in {dt_in} -> {dt} attempts {lw.attempts} red8 {r} owned {d.n_owned} ghost {d.n_ghost}")
        got = by_id([d.local.download() for d in R])
        g = state_arrays(got)
        w = state_arrays(ref[s][0])
        for k in g:
            a, b = np.asarray(g[k], np.float64), np.asarray(w[k], np.float64)
            print(f"   {k}: nan {np.isnan(a).sum()} maxabs {np.nanmax(np.abs(a)):.6g} ref {np.nanmax(np.abs(b)):.6g} diff {np.nanmax(np.abs(a-b)):.3g}")
        print("   ref dt", ref[s][1], "attempts", ref[s][2])
