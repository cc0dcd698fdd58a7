"""In-tree build of the product libraries.

    python -m libclsph_b200.build [--force] [-v]

  libclsph_b200/libsphb.so        the CUDA step engine + its C ABI (include/sphb.h), nvcc, sm_100a.
                                  kernels_exact.cu is compiled with -fmad=false (bit-exact
                                  passes); everything else with nvcc's defaults.
  libclsph_b200/libclsph_host.so  the C++11 host classes that keep libclsph's API (host/),
                                  g++ -std=c++11, linked against libsphb.so, + include/clsph_host.h
  libclsph_b200/clsphparticles    the command-line driver (host/example/particles.cpp)

No torch involvement: plain CUDA runtime + C ABI.
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
# experiments: SPHB_XSUB=<1|2|4> builds the engine with that many x sub-cells per cell of side h
# (csrc/common.cuh: kXSub) into SPHB_OUT (default: the product library)
# SPHB_DEFS="-DNAME=value ..." does the same for any other compile-time parameter
if os.environ.get("SPHB_DEFS"):
    COMMON += os.environ["SPHB_DEFS"].split()
    OBJDIR = os.path.join(OBJDIR, "defs_" + "".join(ch if ch.isalnum() else "_" for ch in os.environ["SPHB_DEFS"]))
    OUT = os.environ.get("SPHB_OUT", os.path.join(OBJDIR, "libsphb.so"))
if os.environ.get("SPHB_XSUB"):
    COMMON.append("-DSPHB_XSUB=" + os.environ["SPHB_XSUB"])
    OBJDIR = os.path.join(OBJDIR, "xsub" + os.environ["SPHB_XSUB"])
    OUT = os.environ.get("SPHB_OUT", os.path.join(OBJDIR, "libsphb.so"))

# source -> extra flags
SOURCES = {
    "engine.cu": [],
    "kernels_exact.cu": ["-fmad=false"],
    "kernels_sort.cu": [],
    "kernels_neighbors.cu": [],
    "kernels_slab.cu": ["-fmad=false"],
    "slab.cu": [],
    "slab_nccl.cu": [],
    "noalloc.cu": ["-fmad=false"],
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


HOST = os.path.join(HERE, "host")
HOST_OUT = os.path.join(HERE, "libclsph_host.so")
CLI_OUT = os.path.join(HERE, "clsphparticles")
HOST_SOURCES = ["scene.cpp", "sph_simulation.cpp", "sph_simulation_slabs.cpp", "capi.cpp", "util/obj_reader.cpp",
                "file_save_delegates/houdini_file_saver.cpp"]
# -ffp-contract=off: scene preparation is compared bit for bit with the reference's
HOST_FLAGS = ["-std=c++11", "-O2", "-ffp-contract=off", "-fPIC", "-Wall", "-pthread",
              "-I", HOST, "-I", os.path.join(HERE, "..", "include")]


def build_host(force=False, verbose=False):
    """libclsph_host.so + clsphparticles (needs libsphb.so to link against)."""
    gxx = shutil.which("g++") or "/usr/bin/g++"
    os.makedirs(os.path.join(OBJDIR, "host"), exist_ok=True)
    hdrs = []
    for root, _, files in os.walk(HOST):
        hdrs += [os.path.join(root, f) for f in files if f.endswith(".h")]
    hdrs += [os.path.join(HERE, "..", "include", f) for f in ("sphb.h", "clsph_host.h")]
    hdrs.append(os.path.abspath(__file__))
    objs = []
    for src in HOST_SOURCES:
        path = os.path.join(HOST, src)
        obj = os.path.join(OBJDIR, "host", src.replace("/", "_").replace(".cpp", ".o"))
        objs.append(obj)
        if force or _stale(obj, [path] + hdrs):
            cmd = [gxx] + HOST_FLAGS + ["-c", path, "-o", obj]
            if verbose:
                print(" ".join(cmd), flush=True)
            subprocess.check_call(cmd)
    link = ["-L", HERE, "-lsphb", "-Wl,-rpath,$ORIGIN", "-pthread"]
    if force or _stale(HOST_OUT, objs + [OUT]):
        cmd = [gxx, "-shared", "-o", HOST_OUT] + objs + link
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    cli_src = os.path.join(HOST, "example", "particles.cpp")
    if force or _stale(CLI_OUT, [cli_src, HOST_OUT] + hdrs):
        cmd = [gxx] + HOST_FLAGS + [cli_src, "-o", CLI_OUT, "-L", HERE, "-lclsph_host", "-lsphb",
                                    "-Wl,-rpath,$ORIGIN", "-pthread"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return HOST_OUT


REFERENCE = "/root/reference"
REF_CALLER_OUT = os.path.join(HERE, "build", "reference_caller", "particles")


def build_reference_caller(force=False, verbose=False):
    """The REFERENCE's own command-line caller — /root/reference/example/particles.cpp, compiled
    where it lies, unchanged — against this repository's headers (host/, include/) and libraries:
    the check that the C++ surface of libclsph/sph_simulation.h:10-27 is kept.
    Only cereal's headers come from the reference tree (-I /root/reference finds util/cereal; every
    other include resolves in host/ first).  The binary lands under build/ (git-ignored; it travels
    to the GPU box, where the reference tree does not exist).  Returns None without the tree."""
    src = os.path.join(REFERENCE, "example", "particles.cpp")
    if not os.path.exists(src):
        return REF_CALLER_OUT if os.path.exists(REF_CALLER_OUT) else None
    gxx = shutil.which("g++") or "/usr/bin/g++"
    os.makedirs(os.path.dirname(REF_CALLER_OUT), exist_ok=True)
    if force or _stale(REF_CALLER_OUT, [src, HOST_OUT, OUT]):
        cmd = [gxx, "-std=c++11", "-O2", "-I", HOST, "-I", os.path.join(HERE, "..", "include"), "-I", REFERENCE,
               src, "-o", REF_CALLER_OUT, "-L", HERE, "-lclsph_host", "-lsphb", "-Wl,-rpath,$ORIGIN/../..",
               "-pthread"]
        if verbose:
            print(" ".join(cmd), flush=True)
        subprocess.check_call(cmd)
    return REF_CALLER_OUT


def build_all(force=False, verbose=False):
    out = build(force, verbose), build_host(force, verbose)
    build_reference_caller(force, verbose)
    return out


if __name__ == "__main__":
    print(build_all(force="--force" in sys.argv, verbose="-v" in sys.argv))
