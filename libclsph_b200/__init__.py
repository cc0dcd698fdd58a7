"""libclsph_b200 — B200-native SPH step engine behind libclsph's sph_simulation API.

The product is the C-ABI library `libsphb.so` (include/sphb.h, built from csrc/ by
libclsph_b200.build) and the C++11 host classes in host/.  This Python package is a thin
ctypes binding of the same C ABI, used by tests/ and bench.py; it has no CPU path: loading
fails loudly if the CUDA library is missing, and sphb_create() fails without a GPU.
"""
from .engine import (  # noqa: F401
    BB,
    PARAMS,
    PARTICLE,
    TERMS,
    PASS_NAMES,
    Engine,
    SphbError,
    lib,
    lib_path,
)
