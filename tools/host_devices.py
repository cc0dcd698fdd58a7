"""sph_simulation::devices end to end: `steps` sub-steps of an n-particle water lattice over G GPUs
through the C++ frame loop (host threads + NCCL), timed against the single-GPU loop.
    python tools/host_devices.py G n steps"""
import os, sys, tempfile, time
import numpy as np
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path[:0] = [ROOT, os.path.join(ROOT, "tests")]
import bench
from libclsph_b200 import host

G, n, steps = int(sys.argv[1]), int(sys.argv[2]), int(sys.argv[3])
with tempfile.TemporaryDirectory() as tmp:
    case = bench.build_case(tmp, n, "water_box")
    case["fluid_json"], case["sim_json"] = os.path.join(tmp, "fluid.json"), os.path.join(tmp, "sim.json")
    case["obj"] = os.path.join(tmp, f"water_box_{n}.obj")
    out = {}
    for g in (1, G):
        t0 = time.perf_counter()
        r = host.simulate(case["fluid_json"], case["sim_json"], case["obj"], device=0, max_steps=steps, n=n, devices=g)
        out[g] = (time.perf_counter() - t0, r)
        print(f"devices {g}: {steps} steps + set-up + 2 frame read-backs in {out[g][0]:.2f} s, "
              f"steps {r['steps']} frames {r['frames']} recuts {r.get('recuts')}", flush=True)
    a, b = out[1][1]["state"], out[G][1]["state"]
    ka, kb = np.sort(a["grid_index"]), np.sort(b["grid_index"])
    print("same key multiset fraction", float(np.mean(ka == kb)))
    for f in ("position", "velocity", "density"):
        print(f, "range", float(np.nanmin(a[f])), float(np.nanmax(a[f])), "|", float(np.nanmin(b[f])), float(np.nanmax(b[f])))
