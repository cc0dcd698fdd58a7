"""debug: slab engine world-1: continuous run vs download + re-upload after every step"""
import sys, os, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from common import make_case, state_arrays
from test_gpu_slab import slab_world, by_id

with tempfile.TemporaryDirectory() as tmp:
    case = make_case(tmp, "water", 32768)
    a = slab_world(case, 1)
    b = slab_world(case, 1)
    dta = dtb = 1.0 / 600.0
    for s in range(4):
        dta = a.step(dta)
        dtb = b.step(dtb)
        ga = [d.local.download() for d in a.ranks]
        gb = [d.local.download() for d in b.ranks]
        sa, sb = state_arrays(by_id(ga)), state_arrays(by_id(gb))
        print(f"step {s} dt {dta} {dtb}")
        for k in sa:
            x, y = np.asarray(sa[k], np.float64), np.asarray(sb[k], np.float64)
            print(f"   {k}: maxabs {np.nanmax(np.abs(x)):.6g} diff {np.nanmax(np.abs(x - y)):.3g} nan {np.isnan(y).sum()}")
        # re-upload b from its own download
        b.start([(gb[0][0].copy(), gb[0][1].copy())])
