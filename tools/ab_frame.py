"""A/B of the stepping modes on identical trajectories: every leg re-uploads the lattice, runs
`warmup` steps and times `steps` steps through sphb_run_frame (CUDA events on the engine's
stream).  python tools/ab_frame.py [n ...]"""
import json
import os
import sys
import tempfile

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import bench  # noqa: E402


def main():
    from libclsph_b200 import Engine, host, lib
    sizes = [int(a) for a in sys.argv[1:]] or [1000000, 16777216]
    steps, warmup, repeats = 10, 3, 3
    legs = [("host", dict(device_loop=0, side_stream=1)), ("host_noside", dict(device_loop=0, side_stream=0)),
            ("queue", dict(device_loop=1, graph=0, side_stream=1)),
            ("graph", dict(device_loop=1, graph=1, side_stream=1)),
            ("graph_noside", dict(device_loop=1, graph=1, side_stream=0)),
            ("host_scan3", dict(device_loop=0, side_stream=1, scan=0)),
            ("graph_scan3", dict(device_loop=1, graph=1, side_stream=1, scan=0))]
    legs = [(k, dict(dict(scan=1), **v)) for k, v in legs]  # (scan: process-wide, so every leg sets it)
    if os.environ.get("AB_LEGS"):
        legs = [(k, v) for k, v in legs if k in os.environ["AB_LEGS"].split(",")]
    out = {}
    with tempfile.TemporaryDirectory() as tmp:
        for n in sizes:
            case = bench.build_case(tmp, n, "dam_break" if n <= 2000000 else "water_box")
            sc = case["scene"]
            stage, stage_ptr = bench.pinned_particles(n)
            host.lattice(n, case["volume"], out=stage)
            e = Engine(case["params"], case["terms"], 0)
            e.set_scene(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
            res = {}
            for rep in range(repeats):
                for name, opts in legs:
                    for k, v in opts.items():
                        e.set_option(k, v)
                    e.upload_ptr(stage_ptr.value, n)
                    dt, _, _ = e.run_frame(1.0 / 600.0, 1.0, warmup)
                    e.event_record(0)
                    dt, _, done = e.run_frame(dt, 1.0, steps)
                    e.event_record(1)
                    res.setdefault(name, []).append(round(e.event_elapsed_ms(0, 1) / done, 4))
                    if rep == 0 and os.environ.get("AB_PASSES"):
                        e.set_option("side_stream", 0)
                        e.profile(True)
                        for _ in range(steps):
                            dt = e.step(dt)
                        ms, _, k = e.pass_times()
                        e.profile(False)
                        e.set_option("side_stream", 1)
                        res[name + "_passes"] = {p: round(v / k, 4) for p, v in ms.items() if v}
            e.close()
            lib().sphb_host_free(stage_ptr)
            out[n] = res
            print(n, json.dumps(res), flush=True)
    return out


if __name__ == "__main__":
    main()
