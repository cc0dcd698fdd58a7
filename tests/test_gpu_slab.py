"""GPU: the slab (multi-GPU) path of the engine.

On ONE GPU: (a) a world_size-1 slab engine against the whole-domain engine, (b) 2 and 3 slab
engines side by side on the same device (LoopbackWorld: the real CUDA pack / unpack / ghost /
commit kernels, messages moved by a device copy) against the whole-domain engine.  With 2+
GPUs: the same comparison with one process per GPU over NCCL (tests/multi_gpu_check.py under
torchrun).

Bar: from an identical state, one step agrees per particle id within TOL = 1e-5 (the ghost layer
changes only the summation order) and grid_index is bit-exact; 4 free-running steps within
50*TOL.
"""
import os
import subprocess
import sys

import numpy as np
import pytest

from common import ROOT, make_case, new_engine, rel_err, state_arrays
from test_gpu_parity import TOL

pytestmark = pytest.mark.gpu


def by_id(parts):
    p = np.concatenate([x[0] for x in parts])
    ids = np.concatenate([x[1] for x in parts])
    assert len(np.unique(ids)) == len(ids)
    return p[np.argsort(ids)]


def reference_run(case, steps):
    """whole-domain engine; returns per step the state in upload (= id) order"""
    out = []
    with new_engine(case) as e:
        e.upload(case["particles"])
        dt = 1.0 / 600.0
        n = len(case["particles"])
        order = np.arange(n)
        for _ in range(steps):
            dt = e.step(dt)
            st = e.download()
            perm = e.debug()["perm"]
            order = order[perm]          # id of the particle now at each sorted position
            inv = np.empty(n, np.int64)
            inv[order] = np.arange(n)
            out.append((st[inv], dt, e.debug()["attempts"]))
    return out


def compare(got, want, tol, where):
    g, w = state_arrays(got), state_arrays(want)
    for k in ("pos", "vel", "vhf", "acc", "rho"):
        err = rel_err(g[k], w[k])
        assert err <= tol, f"{where}: {k} {err:.2e}"


def slab_world(case, world):
    from libclsph_b200.slab import CudaSlab, LoopbackWorld, equal_count_cuts
    p = case["particles"]
    n = len(p)
    ids = np.arange(n, dtype=np.uint32)
    x = p["position"][:, 0]
    cuts = equal_count_cuts(x, world)
    locs = [CudaSlab(case["params"], case["terms"], case["scene"], 0, n, n) for _ in range(world)]
    lw = LoopbackWorld(locs, cuts)
    parts = []
    for r in range(world):
        m = (x >= np.float32(cuts[r])) & (x < np.float32(cuts[r + 1]))
        parts.append((p[m], ids[m]))
    lw.start(parts)
    return lw


def test_inner_slab_narrower_than_the_halo_is_refused(tmp_path):
    """Ranks exchange with their two neighbours only, so an inner slab thinner than the 2 h ghost
    layer (+ h of motion) would silently lose ghosts: sphb_slab_configure refuses it."""
    from libclsph_b200 import SphbError
    from libclsph_b200.slab import CudaSlab
    case = make_case(tmp_path, "water", 4096)
    h = float(case["params"]["h"][0])
    loc = CudaSlab(case["params"], case["terms"], case["scene"], 0, 4096, 4096)
    try:
        with pytest.raises(SphbError, match="inner slab"):
            loc.configure(0.0, 2.9 * h, True, True)
        loc.configure(0.0, 3.1 * h, True, True)    # wide enough
        loc.configure(0.0, 0.5 * h, True, False)   # an edge slab may be thin: its one neighbour owns the rest
    finally:
        loc.close()


@pytest.mark.parametrize("world", [1, 2, 3])
def test_slabs_on_one_gpu_match_whole_domain(tmp_path, world):
    case = make_case(tmp_path, "water", 32768)
    steps = 4
    ref = reference_run(case, steps)
    lw = slab_world(case, world)
    dt = 1.0 / 600.0
    for s in range(steps):
        dt = lw.step(dt)
        got = by_id([d.local.download() for d in lw.ranks])
        want, want_dt, want_attempts = ref[s]
        tol = TOL if s == 0 else 50 * TOL
        compare(got, want, tol, f"world {world} step {s}")
        assert abs(dt - want_dt) <= tol * want_dt
        if s == 0:
            assert np.array_equal(got["grid_index"], want["grid_index"])
            assert lw.attempts == want_attempts
        if world > 1:
            assert all(d.n_ghost > 0 for d in lw.ranks)
    for d in lw.ranks:
        d.local.close()


def test_slab_reupload_between_steps(tmp_path):
    """A slab engine whose state is downloaded and uploaded again after every step (what a hook
    that returns true, or a re-cut of the slabs, does) against one that runs on: the same bits;
    and a download right after an upload returns what was uploaded, derived fields included."""
    case = make_case(tmp_path, "water", 32768)
    a, b = slab_world(case, 1), slab_world(case, 1)
    dta = dtb = 1.0 / 600.0
    for s in range(3):
        dta, dtb = a.step(dta), b.step(dtb)
        assert dta == dtb
        ga, gb = a.ranks[0].local.download(), b.ranks[0].local.download()
        sa, sb = by_id([ga]), by_id([gb])
        assert sa.tobytes() == sb.tobytes(), f"step {s}"
        b.start([(gb[0].copy(), gb[1].copy())])
        back = b.ranks[0].local.download()
        assert np.array_equal(back[1], gb[1]) and back[0].tobytes() == gb[0].tobytes(), f"echo {s}"
    for w in (a, b):
        w.ranks[0].local.close()


def test_slab_capacity_errors(tmp_path):
    from libclsph_b200 import SphbError
    from libclsph_b200.slab import CudaSlab, LoopbackWorld, equal_count_cuts
    case = make_case(tmp_path, "water", 4096)
    p = case["particles"]
    cuts = equal_count_cuts(p["position"][:, 0], 2)
    locs = [CudaSlab(case["params"], case["terms"], case["scene"], 0, 4096, 8) for _ in range(2)]
    lw = LoopbackWorld(locs, cuts)
    x = p["position"][:, 0]
    ids = np.arange(len(p), dtype=np.uint32)
    lw.start([(p[x < np.float32(cuts[1])], ids[x < np.float32(cuts[1])]),
              (p[x >= np.float32(cuts[1])], ids[x >= np.float32(cuts[1])])])
    with pytest.raises(SphbError) as ei:
        lw.step(1.0 / 600.0)
    assert ei.value.code == -6  # SPHB_ERR_CAPACITY: the ghost layer does not fit 8 records


@pytest.mark.parametrize("driver", ["native-p2p", "native-nccl", "python"])
def test_two_processes_over_nccl(tmp_path, driver):
    """One process per GPU: the library's own driver (sphb_slab_step) with the halo written straight
    into the neighbour's inbox over peer memory (CUDA IPC) and with ncclSend / ncclRecv, and the
    Python protocol driver over torch.distributed, each against the whole-domain engine."""
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    port = {"native-p2p": "29516", "native-nccl": "29517", "python": "29518"}[driver]
    extra = {"native-p2p": ["--p2p", "1", "--expect-p2p", "1"], "native-nccl": ["--p2p", "0", "--expect-p2p", "0"],
             "python": []}[driver]
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", port,
           os.path.join(ROOT, "tests", "multi_gpu_check.py"), "--particles", "262144", "--steps", "3",
           "--driver", driver.split("-")[0]] + extra
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MULTI_GPU_CHECK OK" in r.stdout


def test_halo_message_resend_over_nccl(tmp_path):
    """Native driver with no spare room in the trimmed halo messages: a message that grew since the
    previous step is exchanged again at full capacity, and the result is unchanged."""
    import re
    import torch
    if torch.cuda.device_count() < 2:
        pytest.skip("needs 2 GPUs")
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", "--nproc-per-node=2",
           "--master-addr", "127.0.0.1", "--master-port", "29519",
           os.path.join(ROOT, "tests", "multi_gpu_check.py"), "--particles", "262144", "--steps", "6",
           "--driver", "native", "--halo-margin", "0", "--p2p", "0"]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=900)
    assert r.returncode == 0, r.stdout[-3000:] + r.stderr[-3000:]
    assert "MULTI_GPU_CHECK OK" in r.stdout
    resends = [int(m) for m in re.findall(r"halo resends (\d+)", r.stdout)]
    assert len(resends) == 2 and resends[0] == resends[1], r.stdout[-2000:]
    print("resends per rank:", resends)
