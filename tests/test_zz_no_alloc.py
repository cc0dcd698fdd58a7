"""sphb_cell_table_noalloc (include/sphb.h "entry points that allocate nothing") against the
existing entry point for the same product, sphb_download_debug's `cell_table`.

The file is named so that pytest collects it LAST.  The entry
point has no counterpart in the reference: it was added for
callers that cannot call cudaMalloc.  Integer work: the bar is BIT-EXACT, as for the cell table
everywhere else in this suite.  Device memory for the caller's inputs / outputs comes from torch
(the caller may allocate; the entry point may not)."""
import ctypes as C

import numpy as np
import pytest

from common import make_case, new_engine

pytestmark = pytest.mark.gpu


def _call(lib, pos_dev, n, params, table_dev):
    rc = lib.sphb_cell_table_noalloc(C.c_void_p(pos_dev.data_ptr()), n, params.ctypes.data_as(C.c_void_p),
                                     C.c_void_p(table_dev.data_ptr()), table_dev.numel(), None)
    assert rc == 0, (rc, lib.sphb_last_error(None))


def test_cell_table_noalloc_matches_the_engine_twice_in_a_row(tmp_path):
    import torch
    from libclsph_b200 import lib
    L = lib()
    L.sphb_last_error.restype = C.c_char_p
    case = make_case(tmp_path, "water", 4096)
    states, tables, params = [], [], []
    with new_engine(case) as e:
        e.upload(case["particles"])
        dt = 1.0 / 600.0
        for _ in range(2):  # two different states: the lattice, and the lattice one step later
            states.append(e.download())
            dt = e.step(dt)
            tables.append(e.debug()["cell_table"].copy())
            params.append(e.parameters().copy())
    assert not np.array_equal(states[0]["position"], states[1]["position"])
    torch.cuda.set_device(0)
    out = []
    for k in range(2):  # two calls in a row, sharing the static scratch
        n = len(states[k])
        pos = torch.from_numpy(np.ascontiguousarray(states[k]["position"])).cuda()  # n x 4 floats
        ncell = int(params[k]["grid_cell_count"][0])
        assert ncell == len(tables[k]) and ncell > 2048  # several scan tiles: the scratch is used
        table = torch.full((ncell + 7,), 0x5A5A5A5A, dtype=torch.int32, device="cuda")  # stale contents
        _call(L, pos, n, params[k], table)
        torch.cuda.synchronize()
        got = table.cpu().numpy().view(np.uint32)
        assert np.array_equal(got[:ncell], tables[k]), f"call {k}: cell table differs from sphb_download_debug's"
        assert (got[ncell:] == 0x5A5A5A5A).all(), "wrote past grid_cell_count"
        assert got[0] == 0 and got[ncell - 1] <= n
        out.append(got[:ncell].copy())
    # the first state again: the result must not depend on what the previous call left behind
    pos = torch.from_numpy(np.ascontiguousarray(states[0]["position"])).cuda()
    table = torch.zeros(len(tables[0]), dtype=torch.int32, device="cuda")
    _call(L, pos, len(states[0]), params[0], table)
    torch.cuda.synchronize()
    assert np.array_equal(table.cpu().numpy().view(np.uint32), out[0])


def test_cell_table_noalloc_rejects_what_it_cannot_hold(tmp_path):
    import torch
    from libclsph_b200 import lib
    L = lib()
    L.sphb_last_error.restype = C.c_char_p
    case = make_case(tmp_path, "water", 4096)
    with new_engine(case) as e:
        e.upload(case["particles"])
        e.step(1.0 / 600.0)
        params = e.parameters().copy()
    pos = torch.from_numpy(np.ascontiguousarray(case["particles"]["position"])).cuda()
    table = torch.zeros(16, dtype=torch.int32, device="cuda")
    args = (C.c_void_p(pos.data_ptr()), len(case["particles"]), params.ctypes.data_as(C.c_void_p),
            C.c_void_p(table.data_ptr()))
    assert L.sphb_cell_table_noalloc(*args, 16, None) == -6  # SPHB_ERR_CAPACITY: output too small
    big = params.copy()
    big["grid_cell_count"] = 4194304 + 1  # SPHB_NOALLOC_MAX_CELLS + 1
    rc = L.sphb_cell_table_noalloc(args[0], args[1], big.ctypes.data_as(C.c_void_p), args[3], 1 << 30, None)
    assert rc == -6 and b"SPHB_NOALLOC_MAX_CELLS" in L.sphb_last_error(None)
    assert L.sphb_cell_table_noalloc(None, 0, params.ctypes.data_as(C.c_void_p), args[3], 16, None) == -1
    torch.cuda.synchronize()
