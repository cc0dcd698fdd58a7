"""bench.py's N > 1 arm: the slab-decomposed step, one process per GPU over NCCL.

Weak scaling (default): every rank owns one reference lattice block of `--particles` particles
(the N = 1 workload); the blocks stand side by side along x in one open-top box, so the whole
job is an N-times longer water box of N * particles particles.  `--scaling strong` instead
cuts the single N = 1 workload (BASELINE.json configs[3]: the 16.7 M box) into N slabs.
Timing: barrier + synchronize on both sides, CUDA events on each engine's stream, MAX over
ranks; `value` = all particles of all ranks * steps / that time.
"""
import json
import os
import tempfile
import time

import numpy as np


def parity_check(B, dist, rank, world, local, n=262144, steps=3):
    """Correctness BEFORE timing: an n-particle water block stepped `steps` times by the `world`
    slab engines — once with the halo over NVLink peer memory (where it can be mapped), once over
    ncclSend / ncclRecv — against ONE whole-domain engine on rank 0, per particle id.  Bar, as in
    tests/test_gpu_slab.py: step 0 within 1e-5 with bit-exact grid_index and the same number of
    advection attempts, later steps within 5e-4 (free-running trajectories), no particle lost or
    owned twice, the same dt on every rank.  Returns a dict for the JSON line (rank 0) or None."""
    from libclsph_b200 import Engine, host
    from libclsph_b200.slab import CudaSlab, SlabDriver, equal_count_cuts

    def rel(a, b):
        a, b = np.asarray(a, np.float64), np.asarray(b, np.float64)
        return float(np.max(np.abs(a - b)) / max(float(np.max(np.abs(b))), 1e-30))

    fields = dict(pos="position", vel="velocity", vhf="intermediate_velocity", acc="acceleration")
    with tempfile.TemporaryDirectory() as tmp:
        case = B.build_case(tmp, n, "dam_break")
    params, terms, sc = case["params"], case["terms"], case["scene"]
    p = host.lattice(n, case["volume"])
    ids = np.arange(n, dtype=np.uint32)
    ref = []
    if rank == 0:  # the whole-domain engine; states re-ordered to id (= upload) order
        e = Engine(params, terms, local)
        e.set_scene(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
        e.upload(p)
        dt, order = 1.0 / 600.0, np.arange(n)
        for _ in range(steps):
            dt = e.step(dt)
            st, dbg = e.download(), e.debug()
            order = order[dbg["perm"]]
            inv = np.empty(n, np.int64)
            inv[order] = np.arange(n)
            ref.append((st[inv], dt, dbg["attempts"]))
        e.close()
    x = p["position"][:, 0]
    cuts = equal_count_cuts(x, world)
    mine = (x >= np.float32(cuts[rank])) & (x < np.float32(cuts[rank + 1]))
    result = dict(particles=n, steps=steps, tolerance=dict(step0=1e-5, later=5e-4), modes={})
    ok_all = True
    for p2p in (1, 0):
        loc = CudaSlab(params, terms, sc, local, int(1.5 * n / world) + 65536, max(65536, n // 4))
        loc.e.set_option("p2p_halo", p2p)
        loc.comm_init(rank, world, dist)
        drv = SlabDriver(loc, cuts, rank, world, dist)
        drv.start(p[mine], ids[mine])
        dt, worst, ok = 1.0 / 600.0, {}, True
        for s in range(steps):
            dt = drv.step(dt)
            part, pid = loc.download()
            gathered = [None] * world if rank == 0 else None
            dist.gather_object((part, pid, dt, drv.attempts), gathered, dst=0)
            if rank == 0:
                allp = np.concatenate([g[0] for g in gathered])
                allid = np.concatenate([g[1] for g in gathered])
                ok &= len(allid) == n and len(np.unique(allid)) == n
                if not ok:  # (keep taking part in the collectives of the remaining steps)
                    continue
                got = allp[np.argsort(allid)]
                want, want_dt, want_attempts = ref[s]
                tol = 1e-5 if s == 0 else 5e-4
                errs = {k: rel(got[f][:, :3], want[f][:, :3]) for k, f in fields.items()}
                errs["rho"] = rel(got["density"], want["density"])
                errs["dt"] = abs(dt - want_dt) / want_dt
                for k, v in errs.items():
                    worst[k] = max(worst.get(k, 0.0), v)
                ok &= all(v <= tol for v in errs.values()) and all(g[2] == gathered[0][2] for g in gathered)
                if s == 0:
                    ok &= bool(np.array_equal(got["grid_index"], want["grid_index"])) and gathered[0][3] == want_attempts
        name = "peer_memory" if loc.p2p_active() else ("nccl_send_recv" if p2p == 0 else "nccl_send_recv(peer memory not mapped)")
        loc.close()
        dist.barrier()
        if rank == 0:
            result["modes"][name] = dict(ok=bool(ok), worst_relative_error={k: float(f"{v:.3g}") for k, v in worst.items()},
                                         grid_index_step0="bit-exact" if ok else "see ok")
            ok_all &= ok
    flag = [ok_all]
    dist.broadcast_object_list(flag, src=0)
    result["ok"] = bool(flag[0])
    return result


def run_config(a, B, torch, dist, rank, world, local, workload):
    """One timed configuration (a.scaling, a.scene, a.particles, a.steps ...) -> the JSON line's
    dict on rank 0, None elsewhere."""
    from libclsph_b200 import PARTICLE, host
    from libclsph_b200.slab import CudaSlab, SlabDriver, equal_count_cuts
    warmup = max(a.warmup, 3)
    strong = getattr(a, "scaling", "weak") == "strong"
    n0 = int(a.particles)
    n_total = n0 if strong else n0 * world

    with tempfile.TemporaryDirectory() as tmp:
        B.write_json(os.path.join(tmp, "fluid.json"), B.WATER)
        B.write_json(os.path.join(tmp, "sim.json"), dict(B.SIM, particles_count=n0))
        params, terms, vol, _, _ = host.load_settings(os.path.join(tmp, "fluid.json"), os.path.join(tmp, "sim.json"))
        params["particles_count"] = n_total
        params["total_mass"] = np.float32(n_total) * params["particle_mass"]
        side = float(np.cbrt(np.float32(vol)))
        per_side = int(np.ceil(np.cbrt(np.float32(n0))))
        spacing = float(np.float32(side) / np.float32(per_side))
        block = host.lattice(n0, vol)  # the reference lattice of one block
        if strong:
            x = block["position"][:, 0]
            cuts = equal_count_cuts(x, world)
            mine = (x >= np.float32(cuts[rank])) & (x < np.float32(cuts[rank + 1]))
            mine_p = block[mine]
            ids = np.nonzero(mine)[0].astype(np.uint32)
            dims = B.water_box_dims(n0)
        else:
            pitch = np.float32(per_side * spacing)
            block["position"][:, 0] += np.float32(rank) * pitch
            mine_p = block
            ids = (np.arange(n0, dtype=np.uint64) + np.uint64(rank) * np.uint64(n0)).astype(np.uint32)
            x0 = -np.float32(side) / np.float32(2)
            cuts = [-np.inf] + [float(x0 + np.float32(r) * pitch - np.float32(spacing) / 2) for r in range(1, world)] + [np.inf]
            d = B.water_box_dims(n0)
            dims = (d[0], float(d[1] + np.float32(world - 1) * pitch), d[2], d[3], d[4], d[5])
        scene_name = getattr(a, "scene", "box")
        if scene_name != "box":
            if not strong:
                raise SystemExit("--scene needs --scaling strong")
            sc = B.fixture_scene(scene_name + "_water", params)
            workload = (f"synthetic {n0}-particle water lattice in the reference's {scene_name}.obj "
                        f"({sc['faces']} faces, {sc['total_gridpoints']} distance-field points), SURVEY 8(d) config 5")
        else:
            obj = os.path.join(tmp, f"box_rank{rank}.obj")
            B.write_box_obj(obj, *dims)
            sc = host.load_scene(obj, float(params["h"][0]) * 2)
    n_mine = len(mine_p)
    # ghost layer: 2 layers of thickness 2h across the slab's cross-section, with head-room
    planes = 2 * float(params["h"][0]) / spacing + 1  # lattice planes inside a 2h layer
    exch = int(1.5 * per_side * per_side * planes) + 65536
    cap = int(n_mine * 1.25) + 2 * exch
    loc = CudaSlab(params, terms, sc, local, cap, exch)
    if getattr(a, "driver", "native") == "native":
        loc.e.set_option("p2p_halo", int(getattr(a, "p2p", 1)))
        loc.comm_init(rank, world, dist)  # the library drives exchange + all-reduce itself
    drv = SlabDriver(loc, cuts, rank, world, dist)
    drv.start(mine_p, ids)
    del block, mine_p
    e = loc.e
    dt = 1.0 / 600.0
    for _ in range(warmup):
        dt = drv.step(dt)
    l0 = e.launch_count()
    sampler = B.ClockSampler(B.physical_device(local)) if rank == 0 else None
    torch.cuda.synchronize()
    dist.barrier()
    torch.cuda.synchronize()
    if sampler:
        sampler.start()
    bytes0 = drv.bytes_sent
    t0 = time.perf_counter()
    e.event_record(0)
    for _ in range(a.steps):
        dt = drv.step(dt)
    e.event_record(1)
    ms = e.event_elapsed_ms(0, 1)
    torch.cuda.synchronize()
    wall = time.perf_counter() - t0
    bytes_timed = drv.bytes_sent - bytes0
    dist.barrier()
    clocks = sampler.finish() if sampler else None
    launches = e.launch_count() - l0
    # per-pass CUDA-event times and host phase times come from `steps` FURTHER steps: engine
    # profiling synchronises inside the step, which the timed region above must not
    e.profile(True)
    drv.profile = True
    if drv.native:
        loc.native_gpu_segments(reset=True)
    for _ in range(a.steps):
        dt = drv.step(dt)
    segments = loc.native_gpu_segments() if drv.native else {}
    pass_ms, pass_launches, _ = e.pass_times()
    e.profile(False)
    drv.profile = False
    stats = torch.tensor([ms, wall * 1e3, float(launches), float(drv.n_owned), float(drv.n_ghost),
                          float(bytes_timed), drv.phase_s["exchange"] * 1e3, drv.phase_s["forces_advect"] * 1e3,
                          drv.phase_s["retries_commit"] * 1e3] + [pass_ms[k] for k in sorted(pass_ms)]
                         + [segments.get(k, 0.0) for k in CudaSlab.SEGMENTS],
                         dtype=torch.float64, device="cuda")
    allstats = [torch.zeros_like(stats) for _ in range(world)]
    dist.all_gather(allstats, stats)
    owned_total = sum(int(s[3]) for s in allstats)
    # end to end: every rank downloads its slab to the host and uploads it again, every step
    e2e = None
    if a.e2e_steps > 0:
        # host buffers page-locked (sphb_host_alloc), as sph_simulation::simulate() holds its array
        stage, stage_ptr = B.pinned_particles(cap)
        ids_raw, ids_ptr = B.pinned_particles((cap * 4 + PARTICLE.itemsize - 1) // PARTICLE.itemsize)
        ids_stage = ids_raw.view(np.uint8)[: cap * 4].view(np.uint32)
        part, pid = loc.download(stage, ids_stage)  # one warm pass of the host path
        drv.start(part, pid)
        dt = drv.step(dt)
        torch.cuda.synchronize()
        dist.barrier()
        t0 = time.perf_counter()
        for _ in range(a.e2e_steps):
            part, pid = loc.download(stage, ids_stage)
            drv.start(part, pid)
            dt = drv.step(dt)
        torch.cuda.synchronize()
        e2e_s = time.perf_counter() - t0
        t = torch.tensor([e2e_s], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        e2e = dict(value=n_total * a.e2e_steps / float(t[0]), unit="particle-steps/s",
                   h2d_bytes_per_step=84 * n_total, d2h_bytes_per_step=84 * n_total + 32 * world,
                   steps=a.e2e_steps)
    line = None
    if rank == 0:
        ms_max = max(float(s[0]) for s in allstats)
        ms_step = ms_max / a.steps
        peak, peak_src = B.measured_peak()
        names = sorted(pass_ms)
        per_rank = []
        for r, s in enumerate(allstats):
            per_rank.append(dict(rank=r, ms_per_step=float(s[0]) / a.steps, owned=int(s[3]), ghosts=int(s[4]),
                                 halo_bytes_per_step=float(s[5]) / a.steps,
                                 host_phase_ms_per_step=dict(pack_exchange_unpack=float(s[6]) / a.steps,
                                                            sort_to_first_advect=float(s[7]) / a.steps,
                                                            retries_commit=float(s[8]) / a.steps),
                                 pass_ms_per_step={k: float(s[9 + i]) / a.steps for i, k in enumerate(names)},
                                 gpu_segments_ms_per_step={k: float(s[9 + len(names) + i])
                                                           for i, k in enumerate(CudaSlab.SEGMENTS)}))
        slow = max(per_rank, key=lambda d: d["ms_per_step"])
        f_ms = slow["pass_ms_per_step"]["forces"]
        n_loc = slow["owned"]
        achieved = B.ALGO_BYTES["forces"] * n_loc / (f_ms * 1e-3) / 1e9 if f_ms > 0 else 0.0
        line = dict(
            metric="SPH particle-steps/sec", value=n_total * a.steps / (ms_max * 1e-3), unit="particle-steps/s",
            n_gpus=world, steps=a.steps, warmup=warmup, ms_per_step=ms_step, higher_is_better=True,
            scaling="strong" if strong else "weak", vs_baseline=None, dtype="f32", data="synthetic",
            config=dict(workload=workload if strong else
                        f"{world} x ({n0}-particle reference lattice block) side by side along x in one open-top "
                        f"water box (water.json + default.json physics): BASELINE.json configs[3]'s block per GPU",
                        particles=n_total, particles_per_gpu=n0 if not strong else n_total // world,
                        box=[round(float(v), 4) for v in dims],
                        parallelism=f"{world} x-slabs, 2h ghost layer + migration " +
                                    ("written by the pack kernel straight into the neighbour's inbox over NVLink "
                                     "peer memory (CUDA IPC), " if (drv.native and loc.p2p_active()) else
                                     "by NCCL send/recv, ") +
                                    "1 NCCL all-reduce(8 floats) per advection attempt",
                        p2p_halo=bool(drv.native and loc.p2p_active()),
                        driver="sphb_slab_step (C, NCCL by dlopen)" if drv.native else "SlabDriver (Python, torch.distributed)",
                        l2="inputs exceed L2", exchange_capacity_records=exch, owned_total=owned_total),
            clocks=clocks, gpu_launches=int(sum(float(s[2]) for s in allstats)),
            roofline=dict(bound="hbm", kernel="forces pass (k_forces_union), slowest rank", achieved=achieved,
                          peak=peak, unit="GB/s", frac=achieved / peak, traffic=None, peak_source=peak_src,
                          algorithmic_bytes_per_particle=B.ALGO_BYTES["forces"], launch_ms=f_ms),
            ranks=per_rank)
        if e2e:
            line["e2e"] = e2e
    loc.close()
    dist.barrier()
    return line


def main_multi(args, rank, world, local, workload):
    import torch
    import torch.distributed as dist
    import __main__ as B  # bench.py itself (it owns the real stdout)
    if not hasattr(B, "emit"):
        import bench as B
    from libclsph_b200 import PARTICLE, build, host
    from libclsph_b200.slab import CudaSlab, SlabDriver, equal_count_cuts

    torch.cuda.set_device(local)
    dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    if rank == 0:
        build.build_all()
    dist.barrier()
    parity = None
    if not getattr(args, "no_parity", False):
        parity = parity_check(B, dist, rank, world, local)
        if not parity["ok"]:  # a wrong answer is not timed
            if rank == 0:
                B.emit(dict(metric="SPH particle-steps/sec", value=None, n_gpus=world, error="parity_check failed",
                            parity_check=parity))
            dist.destroy_process_group()
            return 1
    import argparse
    line = run_config(args, B, torch, dist, rank, world, local, workload)
    secondary = {}
    if not getattr(args, "no_secondary", False) and getattr(args, "scaling", "weak") == "weak" \
            and getattr(args, "scene", "box") == "box":
        # what BASELINE.json configs[3] / configs[4] word, recorded by the driver's default command:
        # the STRONG series (the 16.7 M box cut into N slabs) and the collision-heavy scenes
        legs = [("strong_16M", dict(scaling="strong", scene="box", particles=args.particles))]
        legs += [(f"strong_{sc}_4M", dict(scaling="strong", scene=sc, particles=4096000)) for sc in ("labyrinth", "river")]
        for name, over in legs:
            a2 = argparse.Namespace(**dict(vars(args), e2e_steps=0, **over))
            l2 = run_config(a2, B, torch, dist, rank, world, local,
                            f"synthetic {over['particles']}-particle water box (BASELINE.json configs[3]) cut into "
                            f"{world} x-slabs")
            if rank == 0:
                secondary[name] = dict(workload=l2["config"]["workload"], scaling="strong", value=l2["value"],
                                       unit=l2["unit"], ms_per_step=l2["ms_per_step"], particles=l2["config"]["particles"],
                                       p2p_halo=l2["config"]["p2p_halo"],
                                       ranks=[dict(rank=r["rank"], ms_per_step=r["ms_per_step"], owned=r["owned"],
                                                   ghosts=r["ghosts"], host_phase_ms_per_step=r["host_phase_ms_per_step"],
                                                   gpu_segments_ms_per_step=r["gpu_segments_ms_per_step"])
                                              for r in l2["ranks"]])
    if rank == 0:
        line["parity_check"] = parity
        line["config"]["secondary"] = secondary or None
        B.emit(line)
    dist.barrier()
    dist.destroy_process_group()
    return 0
