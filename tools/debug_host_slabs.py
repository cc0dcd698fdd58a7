import sys, os, tempfile
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
from common import make_case
from libclsph_b200 import host
n, steps = 32768, int(sys.argv[1]) if len(sys.argv) > 1 else 5
with tempfile.TemporaryDirectory() as tmp:
    case = make_case(tmp, "water", n)
    one = host.simulate(case["fluid_json"], case["sim_json"], case["obj"], max_steps=steps, n=n)
    many = host.simulate(case["fluid_json"], case["sim_json"], case["obj"], max_steps=steps, n=n, devices=-1,
                         rebalance=float(sys.argv[2]) if len(sys.argv) > 2 else 1.25)
    print(one["steps"], many["steps"], one["frames"], many["frames"])
    np.savez_compressed(os.path.join(ROOT, "gpurun_out", f"dbg_host_slabs_{steps}.npz"), one=one["state"], many=many["state"],
                        p1=one["params"], p2=many["params"])
