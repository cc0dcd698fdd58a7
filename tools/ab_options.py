#!/usr/bin/env python
"""A/B of engine options on one box: ms per step of sphb_run_frame over the headline workload.

    python tools/ab_options.py --sets "side_stream=0;side_stream=1;side_stream=0,device_loop=1" [--particles N]
"""
import argparse
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=16777216)
    ap.add_argument("--kind", default="water_box")
    ap.add_argument("--steps", type=int, default=20)
    ap.add_argument("--warmup", type=int, default=5)
    ap.add_argument("--sets", default="side_stream=1;side_stream=0")
    ap.add_argument("--repeat", type=int, default=2)
    a = ap.parse_args()
    from libclsph_b200 import Engine, host
    with tempfile.TemporaryDirectory() as tmp:
        case = bench.build_case(tmp, a.particles, a.kind)
    n, sc = case["n"], case["scene"]
    stage, stage_ptr = bench.pinned_particles(n)
    host.lattice(n, case["volume"], out=stage)
    e = Engine(case["params"], case["terms"], 0)
    e.set_scene(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
    for rep in range(a.repeat):
        for st in a.sets.split(";"):
            opts = [kv.split("=") for kv in st.split(",") if kv]
            for k, v in opts:
                e.set_option(k, int(v))
            e.upload_ptr(stage_ptr.value, n)
            dt, _, _ = e.run_frame(1.0 / 600.0, 1.0, a.warmup)
            e.event_record(0)
            dt, _, done = e.run_frame(dt, 1.0, a.steps)
            e.event_record(1)
            print(json.dumps(dict(options=st, ms_per_step=e.event_elapsed_ms(0, 1) / done, steps=done, rep=rep)), flush=True)
    e.close()


if __name__ == "__main__":
    main()
