#!/usr/bin/env python
"""Houdini .geo writer throughput: this repository's houdini_file_saver against the REFERENCE's own
(libclsph/file_save_delegates/houdini_file_saver.cpp + util/houdini_geo/HoudiniFileDumpHelper.cpp:19-90,
linked into oracle/_ref as orcgeo_write), one frame each to the same file system, same particles.

    python tools/geo_throughput.py [--particles 1000000,16777216] [--out profiles/r02_geo_throughput.json]

Both write <tmp>/frames/frame0000001.geo; the files are compared byte for byte.  MB/s = file size /
wall time of the writeFrameToFile call.  Runs on the CPU (needs oracle/_ref, i.e. the build
container); SURVEY.md 8(f)-2."""
import argparse
import ctypes as C
import filecmp
import json
import os
import sys
import tempfile
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", default="1000000,16777216")
    ap.add_argument("--out", default=os.path.join(ROOT, "profiles", "r02_geo_throughput.json"))
    a = ap.parse_args()
    from common import DEFAULT_SIM, FLUIDS
    from libclsph_b200 import host
    from oracle import pyorc
    orc = pyorc.load("reference")
    orc.lib.orcgeo_write.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_int]
    rows = []
    for n in [int(x) for x in a.particles.split(",")]:
        params, _, vol = orc.settings(FLUIDS["water"], DEFAULT_SIM, particles_count=n)
        p = host.lattice(n, float(vol))
        rng = np.random.default_rng(3)
        p["velocity"][:, :3] = rng.normal(0, 1.5, (n, 3)).astype(np.float32)
        p["density"] = rng.uniform(900, 1100, n).astype(np.float32)
        with tempfile.TemporaryDirectory() as ta, tempfile.TemporaryDirectory() as tb:
            os.makedirs(os.path.join(ta, "frames"))
            os.makedirs(os.path.join(tb, "frames"))
            t0 = time.perf_counter()
            host.save_frame(ta + "/", p, params)
            ours = time.perf_counter() - t0
            t0 = time.perf_counter()
            orc.lib.orcgeo_write((tb + "/").encode(), p.ctypes.data, params.ctypes.data, 1)
            ref = time.perf_counter() - t0
            fa, fb = (os.path.join(t, "frames", "frame0000001.geo") for t in (ta, tb))
            size = os.path.getsize(fa)
            same = filecmp.cmp(fa, fb, shallow=False)
        rows.append(dict(particles=n, file_bytes=size, identical=bool(same),
                         ours=dict(seconds=ours, mb_per_s=size / ours / 1e6),
                         reference=dict(seconds=ref, mb_per_s=size / ref / 1e6), speedup=ref / ours))
        print(json.dumps(rows[-1]), flush=True)
    doc = dict(what=".geo frame writer, this repository against the reference's own writer (oracle/_ref)",
               host_threads=len(os.sched_getaffinity(0)), rows=rows)
    json.dump(doc, open(a.out, "w"), indent=1)


if __name__ == "__main__":
    main()
