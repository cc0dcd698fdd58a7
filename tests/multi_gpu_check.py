"""One process per GPU over NCCL (run under torchrun): the slab-decomposed step against the
whole-domain engine on rank 0, per particle id.

    python -m torch.distributed.run --nnodes=1 --nproc-per-node=2 --master-addr 127.0.0.1 \
        --master-port 29517 tests/multi_gpu_check.py --particles 262144 --steps 3

Prints `MULTI_GPU_CHECK OK` on rank 0 when every step agrees (step 0 within 1e-5 and with
bit-exact grid_index; later steps within 5e-4, trajectories being chaotic)."""
import argparse
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
sys.path.insert(0, HERE)


def main():
    import torch
    import torch.distributed as dist
    from common import make_case, rel_err, state_arrays
    from libclsph_b200.slab import CudaSlab, SlabDriver, equal_count_cuts
    from test_gpu_slab import reference_run

    ap = argparse.ArgumentParser()
    ap.add_argument("--particles", type=int, default=262144)
    ap.add_argument("--steps", type=int, default=3)
    ap.add_argument("--driver", default="native", choices=["native", "python"])
    ap.add_argument("--p2p", type=int, default=1, help="native driver: 1 = halo over peer memory when it can be "
                    "mapped, 0 = ncclSend / ncclRecv")
    ap.add_argument("--expect-p2p", type=int, default=-1, help="fail unless the peer-memory exchange is (1) / is not (0) active")
    ap.add_argument("--halo-margin", type=int, default=-1,
                    help="native driver: spare records per halo message (0 forces the resend path)")
    a = ap.parse_args()
    rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    with tempfile.TemporaryDirectory() as tmp:
        case = make_case(tmp, "water", a.particles)
    p = case["particles"]
    n = len(p)
    ids = np.arange(n, dtype=np.uint32)
    x = p["position"][:, 0]
    cuts = equal_count_cuts(x, world)
    mine = (x >= np.float32(cuts[rank])) & (x < np.float32(cuts[rank + 1]))
    cap = int(1.5 * n / world) + 65536
    loc = CudaSlab(case["params"], case["terms"], case["scene"], local, cap, max(65536, n // 4))
    if a.driver == "native":
        loc.e.set_option("p2p_halo", a.p2p)
        loc.comm_init(rank, world, dist)
        if a.expect_p2p >= 0:
            assert loc.p2p_active() == bool(a.expect_p2p), f"p2p_active = {loc.p2p_active()}"
        if a.halo_margin >= 0:
            loc.e.set_option("halo_margin", a.halo_margin)
    drv = SlabDriver(loc, cuts, rank, world, dist)
    assert drv.native == (a.driver == "native")
    drv.start(p[mine], ids[mine])
    ref = reference_run(case, a.steps) if rank == 0 else None
    dt = 1.0 / 600.0
    ok = True
    for s in range(a.steps):
        dt = drv.step(dt)
        part, pid = loc.download()
        gathered = [None] * world if rank == 0 else None
        dist.gather_object((part, pid, dt, drv.attempts, drv.n_owned, drv.n_ghost), gathered, dst=0)
        if rank == 0:
            allp = np.concatenate([g[0] for g in gathered])
            allid = np.concatenate([g[1] for g in gathered])
            assert len(np.unique(allid)) == n, "a particle is owned twice or lost"
            got = allp[np.argsort(allid)]
            want, want_dt, want_attempts = ref[s]
            tol = 1e-5 if s == 0 else 5e-4
            g, w = state_arrays(got), state_arrays(want)
            errs = {k: rel_err(g[k], w[k]) for k in ("pos", "vel", "vhf", "acc", "rho")}
            same_dt = all(gg[2] == gathered[0][2] for gg in gathered)
            line = f"step {s}: " + " ".join(f"{k} {v:.1e}" for k, v in errs.items()) + \
                   f" dt {dt:.6g} vs {want_dt:.6g} owned {[gg[4] for gg in gathered]} ghosts {[gg[5] for gg in gathered]}"
            print(line, flush=True)
            ok &= all(v <= tol for v in errs.values()) and same_dt and abs(dt - want_dt) <= tol * want_dt
            if s == 0:
                ok &= bool(np.array_equal(got["grid_index"], want["grid_index"]))
                ok &= gathered[0][3] == want_attempts
    dist.barrier()
    if a.driver == "native":
        print(f"rank {rank}: halo resends {loc.native_resends()} bytes sent {loc.native_bytes_sent()} "
              f"p2p {int(loc.p2p_active())}", flush=True)
    if rank == 0:
        print("MULTI_GPU_CHECK OK" if ok else "MULTI_GPU_CHECK FAILED", flush=True)
    loc.close()
    dist.destroy_process_group()
    return 0 if ok or rank != 0 else 1


if __name__ == "__main__":
    sys.exit(main())
