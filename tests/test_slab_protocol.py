"""CPU, world_size 2 and 3 over gloo: the slab exchange protocol of libclsph_b200/slab.py
(SlabDriver) with the numpy stand-in engine of tests/slab_numpy.py.

Checked after every step against a world_size-1 run of the same driver:
  * every particle is owned by exactly one rank, the one whose slab holds its x;
  * positions are bit-identical (the stand-in's motion does not depend on summation order);
  * the neighbour count within h of every owned particle equals the single-rank count, i.e.
    the 2h ghost layer delivered every particle a rank needs;
  * dt and the number of advection attempts are the same on every rank.
"""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

from libclsph_b200.slab import SlabDriver, equal_count_cuts, message_bytes, needs_retry
from slab_numpy import NumpySlab

H = 0.1
STEPS = 6


def make_particles(n=3000, seed=5):
    from libclsph_b200.engine import PARTICLE
    rng = np.random.default_rng(seed)
    p = np.zeros(n, PARTICLE)
    p["position"][:, :3] = rng.uniform([-1.0, 0.0, -0.5], [1.0, 1.0, 0.5], (n, 3)).astype(np.float32)
    v = rng.normal(0, 4.0, (n, 3)).astype(np.float32)
    v[:, 0] += np.where(p["position"][:, 0] < 0, 25.0, -25.0).astype(np.float32)  # streams cross the faces
    p["intermediate_velocity"][:, :3] = v
    p["velocity"][:, :3] = v
    return p


def run_rank(rank, world, port, out):
    if world > 1:
        os.environ["MASTER_ADDR"] = "127.0.0.1"
        os.environ["MASTER_PORT"] = str(port)
        dist.init_process_group("gloo", rank=rank, world_size=world)
    p = make_particles()
    ids = np.arange(len(p), dtype=np.uint32)
    cuts = equal_count_cuts(p["position"][:, 0], world)
    x = p["position"][:, 0]
    mine = (x >= np.float32(cuts[rank])) & (x < np.float32(cuts[rank + 1]))
    drv = SlabDriver(NumpySlab(H, exchange_capacity=len(p)), cuts, rank, world, dist if world > 1 else None)
    drv.start(p[mine], ids[mine])
    dt = 1.0 / 600.0
    log = []
    for _ in range(STEPS):
        dt = drv.step(dt)
        pos, pid, cnt = drv.local.download()
        log.append(dict(dt=dt, attempts=drv.attempts, pos=pos, ids=pid, count=cnt, n_ghost=drv.n_ghost,
                        gmin=drv.gmin.copy(), gmax=drv.gmax.copy(), cuts=cuts))
    out[rank] = log
    if world > 1:
        dist.destroy_process_group()


def free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def launch(world):
    if world == 1:
        out = {}
        run_rank(0, 1, 0, out)
        return out
    mgr = mp.Manager()
    out = mgr.dict()
    mp.spawn(run_rank, args=(world, free_port(), out), nprocs=world, join=True)
    return dict(out)


@pytest.fixture(scope="module")
def single():
    return launch(1)[0]


@pytest.mark.parametrize("world", [2, 3])
def test_slab_protocol_matches_single_rank(single, world):
    logs = launch(world)
    assert sorted(logs) == list(range(world))
    migrated = 0
    for s in range(STEPS):
        ref = single[s]
        order = np.argsort(ref["ids"])
        ids = np.concatenate([logs[r][s]["ids"] for r in range(world)])
        pos = np.concatenate([logs[r][s]["pos"] for r in range(world)])
        cnt = np.concatenate([logs[r][s]["count"] for r in range(world)])
        # every particle exactly once
        assert np.array_equal(np.sort(ids), ref["ids"][order])
        o = np.argsort(ids)
        assert np.array_equal(pos[o].view(np.uint32), ref["pos"][order].view(np.uint32)), f"step {s}"
        assert np.array_equal(cnt[o], ref["count"][order]), f"step {s}: a rank missed neighbours"
        for r in range(world):
            L = logs[r][s]
            assert L["dt"] == ref["dt"] and L["attempts"] == ref["attempts"]
            assert np.array_equal(L["gmin"], ref["gmin"]) and np.array_equal(L["gmax"], ref["gmax"])
            if world > 1:
                assert L["n_ghost"] > 0
        # ownership follows x: the positions a rank holds were inside its slab before this
        # step's advection, so compare with the PREVIOUS step's positions
        if s > 0:
            prev = {int(i): p for i, p in zip(single[s - 1]["ids"], single[s - 1]["pos"])}
            for r in range(world):
                L = logs[r][s]
                lo, hi = np.float32(L["cuts"][r]), np.float32(L["cuts"][r + 1])
                px = np.array([prev[int(i)][0] for i in L["ids"]], np.float32)
                assert np.all((px >= lo) & (px < hi)), f"step {s} rank {r}"
                migrated += len(set(map(int, L["ids"])) - set(map(int, logs[r][s - 1]["ids"])))
    assert migrated > 0, "the test case must exercise migration"


def test_retry_predicate_and_message_size():
    assert message_bytes(10) == 16 + 480
    assert needs_retry(1.0 / 600.0, 1.0 / 600.0 - 2e-5)
    assert not needs_retry(1.0 / 600.0, 1.0 / 600.0 - 5e-6)


def test_equal_count_cuts_never_split_a_lattice_plane():
    x = np.repeat(np.arange(10, dtype=np.float32) * np.float32(0.1), 7)
    cuts = equal_count_cuts(x, 4)
    assert cuts[0] == -np.inf and cuts[-1] == np.inf
    for c in cuts[1:-1]:
        assert not np.any(x == np.float32(c))
    counts = [int(((x >= np.float32(cuts[r])) & (x < np.float32(cuts[r + 1]))).sum()) for r in range(4)]
    assert sum(counts) == len(x) and min(counts) >= 14


def test_loopback_world_matches_single_rank(single):
    """LoopbackWorld (all ranks in one process, messages moved by copy) runs the same protocol."""
    from libclsph_b200.slab import LoopbackWorld
    p = make_particles()
    ids = np.arange(len(p), dtype=np.uint32)
    world = 3
    cuts = equal_count_cuts(p["position"][:, 0], world)
    x = p["position"][:, 0]
    lw = LoopbackWorld([NumpySlab(H, exchange_capacity=len(p)) for _ in range(world)], cuts)
    parts = []
    for r in range(world):
        m = (x >= np.float32(cuts[r])) & (x < np.float32(cuts[r + 1]))
        parts.append((p[m], ids[m]))
    lw.start(parts)
    dt = 1.0 / 600.0
    for s in range(STEPS):
        dt = lw.step(dt)
        ref = single[s]
        got = [d.local.download() for d in lw.ranks]
        gid = np.concatenate([g[1] for g in got])
        gpos = np.concatenate([g[0] for g in got])
        gcnt = np.concatenate([g[2] for g in got])
        o, ro = np.argsort(gid), np.argsort(ref["ids"])
        assert np.array_equal(gid[o], ref["ids"][ro])
        assert np.array_equal(gpos[o].view(np.uint32), ref["pos"][ro].view(np.uint32))
        assert np.array_equal(gcnt[o], ref["count"][ro])
        assert dt == ref["dt"] and lw.attempts == ref["attempts"]
