"""In-tree build of the sphb C-ABI library (libclsph_b200/libsphb.so) with nvcc for sm_100a.

    python -m libclsph_b200.build [--force]

kernels_exact.cu is compiled with -fmad=false (bit-exact passes); everything else with
nvcc's defaults.  No torch involvement: the library is plain CUDA runtime + C ABI.
"""
import os
import shutil
import subprocess
import sys

HERE = os.path.dirname(os.path.abspath(__file__))
CSRC = os.path.join(HERE, "csrc")
OUT = os.path.join(HERE, "libsphb.so")
OBJDIR = os.path.join(HERE, "build")
ARCH = ["-gencode", "arch=compute_100a,code=sm_100a"]
COMMON = ["-O3", "-std=c++17", "-lineinfo", "-Xcompiler", "-fPIC",
          "-DSPHB_BUILD"]

# source -> extra flags
SOURCES = {
    "engine.cu": [],
    "kernels_exact.cu": ["-fmad=false"],
    "kernels_sort.cu": [],
    "kernels_neighbors.cu": [],
    "multi.cu": [],
}


def nvcc():
    exe = shutil.which("nvcc") or "/usr/local/cuda/bin/nvcc"
    if not os.path.exists(exe):
        raise RuntimeError("nvcc not found: the sphb engine has no non-CUDA build")
    return exe


def _deps():
    hdrs = [os.path.join(CSRC, f) for f in os.listdir(CSRC) if f.endswith((".cuh", ".h"))]
    hdrs.append(os.path.join(HERE, "..", "include", "sphb.h"))
    hdrs.append(os.path.abspath(__file__))
    return hdrs


def _stale(target, sources):
    if not os.path.exists(target):
        return True
    t = os.path.getmtime(target)
    return any(os.path.getmtime(s) > t for s in sources if os.path.exists(s))


def build(force=False, verbose=False):
    os.makedirs(OBJDIR, exist_ok=True)
    objs = []
    deps = _deps()
    for src, extra in SOURCES.items():
        path = os.path.join(CSRC, src)
        if not os.path.exists(path):
            continue
        obj = os.path.join(OBJDIR, src.replace(".cu", ".o"))
        objs.append(obj)
        if force or _stale(obj, [path] + deps):
            cmd = [nvcc()] + ARCH + COMMON + extra + (["-Xptxas", "-v"] if verbose else []) + ["-c", path, "-o", obj]
            if verbose:
                print(" ".join(cmd), flush=True)
            subprocess.check_call(cmd)
    if force or _stale(OUT, objs):
        cmd = [nvcc()] + ARCH + ["-shared", "-o", OUT] + objs + ["-ldl"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return OUT


if __name__ == "__main__":
    print(build(force="--force" in sys.argv, verbose="-v" in sys.argv))
