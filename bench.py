#!/usr/bin/env python
"""bench.py — SPH particle-steps/s of the sphb step engine on B200 (BASELINE.json's metric).

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--particles P]

One "step" = one simulate_single_frame (libclsph/sph_simulation.cpp:200-263) over the whole
particle set.  Workload at N = 1: BASELINE.json configs[3] on one GPU — the synthetic 16 777 216
particle water box (256^3 reference lattice, water.json + default.json physics, open-top box
x,z in [-4.8, 4.8], y in [-0.06, 12]); `config.secondary` adds configs[2] (1 M dam break).
Prints ONE JSON line (rank 0).

  value        particle-steps/s, state resident in HBM, CUDA events on the engine's stream
  e2e          same metric through the C ABI with HOST buffers: every step uploads the 80-byte
               AoS particles from pinned memory, steps, and downloads them again (the
               write_all_frames + pre_frame-hook path of sph_simulation::simulate)
  roofline     the dominant kernel (forces pass): algorithmic bytes (44 B/particle)
               / its CUDA-event time, against MEASURED_PEAKS.json's HBM copy bandwidth
  cpu_baseline the reference's own kernel source compiled as C++ (oracle/_ref) on the host
               cores, on a bounded sample of the same workload
--impl reference times only that CPU arm.
"""
import argparse
import json
import os
import sys
import tempfile
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

WATER = dict(fluid_density=998.29, dynamic_viscosity=3.5, restitution=0, k=100,
             surface_tension_threshold=7.065, surface_tension=0.0728,
             particles_inside_influence_radius=20)
SIM = dict(particles_count=64000, particle_mass=0.05, simulation_time=3, target_fps=60,
           simulation_scale=0.1, write_all_frames=False, serialize=False,
           constant_acceleration=dict(x=0, y=-9.8, z=0))
# open-top box, vertex order / winding of scenes/box.obj (outward normals)
BOX_FACES = [(1, 2, 3), (1, 5, 6), (2, 6, 3), (3, 7, 4), (5, 1, 4), (4, 1, 3), (2, 1, 6), (6, 7, 3),
             (7, 8, 4), (8, 5, 4)]
ALGO_BYTES = dict(bbox=12, locate=16, sort=None, reorder=76, celltable=5, density=20, forces=44,
                  advect=72 + 24, refrank=None)  # advect carries the dt + bbox reductions
STEP_BYTES_FUSED = {3: 285, 4: 301}  # per particle-step, by radix passes


def write_box_obj(path, xmin, xmax, ymin, ymax, zmin, zmax):
    v = [(xmax, ymin, zmin), (xmax, ymin, zmax), (xmin, ymin, zmax), (xmin, ymin, zmin),
         (xmax, ymax, zmin), (xmax, ymax, zmax), (xmin, ymax, zmax), (xmin, ymax, zmin)]
    with open(path, "w") as fh:
        fh.write("o box\n")
        for p in v:
            fh.write("v %.9g %.9g %.9g\n" % p)
        for f in BOX_FACES:
            fh.write("f %d %d %d\n" % f)


def lattice_side(n):
    return float(np.cbrt(np.float32(n) * np.float32(SIM["particle_mass"]) / np.float32(WATER["fluid_density"])))


def water_box_dims(n):
    """configs[3]: the lattice block fills the box wall to wall (16.7 M: x,z in [-4.8, 4.8])."""
    side = lattice_side(n)
    half = side / 2 + 0.082  # 0.082 m clearance, as at 16.7 M
    return (-half, half, -0.06, 1.27 * side, -half, half)


def dam_break_dims(n):
    """configs[2]: block against the -x wall, free to collapse along +x (1 M: x in [-1.9, 5.6])."""
    side = lattice_side(n)
    half = side / 2 + 0.057
    return (-half, half + side, -0.06, 2.17 * side, -half, half)


def write_json(path, d):
    with open(path, "w") as fh:
        fh.write(json.dumps(d, indent=4) + ";")


MUCUS = dict(fluid_density=1000, dynamic_viscosity=36, restitution=0.5, k=5, surface_tension_threshold=5,
             surface_tension=6, particles_inside_influence_radius=40)


def fixture_scene(name, params):
    z = np.load(os.path.join(ROOT, "tests", "golden", f"scene_{name}.npz"))
    assert np.float32(z["threshold"]) == np.float32(params["h"][0]) * 2
    return dict(bbs=z["bbs"], transforms=z["transforms"], rvertices=z["rvertices"],
                total_gridpoints=int(z["total_gridpoints"]), faces=len(z["bbs"]))


def build_case(tmp, n, kind):
    """Settings / lattice / scene through the PRODUCT's host layer (libclsph_b200.host)."""
    from libclsph_b200 import host
    write_json(os.path.join(tmp, "fluid.json"), MUCUS if kind == "mucus_monkeybox" else WATER)
    write_json(os.path.join(tmp, "sim.json"), dict(SIM, particles_count=int(n)))
    params, terms, vol, _, _ = host.load_settings(os.path.join(tmp, "fluid.json"), os.path.join(tmp, "sim.json"))
    if kind == "mucus_monkeybox" or kind.startswith("scene:"):
        # a reference scene (monkeybox.obj: 978 faces, labyrinth.obj: 606, river.obj: 332 ...) as
        # the reference's scene::load made it (tests/golden/make_golden.py); the .obj files
        # themselves are reference data and are not in the repo
        name = "monkeybox_mucus" if kind == "mucus_monkeybox" else kind[6:] + "_water"
        sc = fixture_scene(name, params)
        return dict(params=params, terms=terms, volume=vol, scene=sc, dims=None, n=int(n))
    dims = water_box_dims(n) if kind == "water_box" else dam_break_dims(n)
    obj = os.path.join(tmp, f"{kind}_{n}.obj")
    write_box_obj(obj, *dims)
    sc = host.load_scene(obj, float(params["h"][0]) * 2)
    if sc is None:
        raise RuntimeError("scene load failed")
    return dict(params=params, terms=terms, volume=vol, scene=sc, dims=dims, n=int(n))


class ClockSampler(threading.Thread):
    """Samples SM clock and throttle reasons of one GPU through NVML while the timed region runs."""

    REASONS = {0x8: "hw_slowdown", 0x40: "hw_thermal_slowdown", 0x20: "sw_thermal_slowdown",
               0x4: "sw_power_cap", 0x80: "hw_power_brake_slowdown"}

    def __init__(self, index):
        super().__init__(daemon=True)
        self.index, self.samples, self.reasons, self.max_mhz = index, [], set(), None
        self._stop_evt = threading.Event()
        try:
            import pynvml
            pynvml.nvmlInit()
            self.nv = pynvml
            self.h = pynvml.nvmlDeviceGetHandleByIndex(index)
            self.max_mhz = pynvml.nvmlDeviceGetMaxClockInfo(self.h, pynvml.NVML_CLOCK_SM)
        except Exception:
            self.nv = None

    def run(self):
        if self.nv is None:
            return
        while not self._stop_evt.is_set():
            try:
                self.samples.append(self.nv.nvmlDeviceGetClockInfo(self.h, self.nv.NVML_CLOCK_SM))
                r = self.nv.nvmlDeviceGetCurrentClocksEventReasons(self.h)
                for bit, name in self.REASONS.items():
                    if r & bit:
                        self.reasons.add(name)
            except Exception:
                pass
            self._stop_evt.wait(0.005)

    def finish(self):
        self._stop_evt.set()
        self.join(timeout=2)
        return dict(sm_mhz=float(np.median(self.samples)) if self.samples else None,
                    sm_max_mhz=self.max_mhz, reasons=sorted(self.reasons), samples=len(self.samples))


def physical_device(local_index):
    vis = os.environ.get("CUDA_VISIBLE_DEVICES")
    if vis:
        try:
            return int(vis.split(",")[local_index])
        except Exception:
            return local_index
    return local_index


def kernel_source_hash():
    """sha256 (16 hex digits) of the sources the two neighbour passes are built from: what an ncu
    capture under profiles/ was taken on, and what this run executes."""
    import hashlib
    h = hashlib.sha256()
    for f in ("kernels_neighbors.cu", "neighbors_common.cuh", "neighbors_v1.cuh", "common.cuh", "kernels_exact.cu",
              "kernels_sort.cu"):
        h.update(open(os.path.join(ROOT, "libclsph_b200", "csrc", f), "rb").read())
    return h.hexdigest()[:16]


def ncu_counters(which, n):
    """(record, reason): the committed ncu counters of neighbour pass `which` at n particles, or
    (None, why not) when there is none for the sources this run executes."""
    path = os.path.join(ROOT, "profiles", "ncu_traffic.json")
    try:
        doc = json.load(open(path))
    except Exception as ex:
        return None, f"profiles/ncu_traffic.json unreadable ({ex})"
    if doc.get("source_hash") != kernel_source_hash():
        return None, (f"stale: profiles/ncu_traffic.json was captured on kernel sources {doc.get('source_hash')} "
                      f"(commit {doc.get('commit')}), this run executes {kernel_source_hash()}")
    rec = doc.get(which, {}).get(str(n))
    return (rec, None) if rec else (None, f"no capture at {n} particles")


def opencl_cpu_runtime():
    """BASELINE.md section 1 / SURVEY 8(d): probe for an OpenCL CPU runtime the reference's own host
    path could run on (it is timed beside the GPU only if one exists)."""
    import glob
    found = []
    for pat in ("/etc/OpenCL/vendors/*.icd", "/usr/lib/x86_64-linux-gnu/libpocl*", "/usr/lib*/libpocl*",
                "/usr/local/lib/libpocl*", "/usr/include/CL/cl.h", "/usr/lib/x86_64-linux-gnu/libOpenCL.so*"):
        found += glob.glob(pat)
    cpu = [f for f in found if "pocl" in f.lower() or "intel" in f.lower()]
    if not cpu:
        return "none found" + (f" (only: {', '.join(sorted(found))})" if found else "")
    return "found: " + ", ".join(sorted(found))


def measured_peak():
    path = os.path.join(ROOT, "MEASURED_PEAKS.json")
    if os.path.exists(path):
        try:
            return float(json.load(open(path))["hbm_gbs"]), "measured (MEASURED_PEAKS.json hbm_gbs)"
        except Exception:
            pass
    return 6650.0, "fallback (B200_PROFILING.md 6.65 TB/s)"


def pinned_particles(n):
    """n 80-byte particles in page-locked host memory (sphb_host_alloc) as a numpy view."""
    import ctypes as C
    from libclsph_b200 import PARTICLE, lib
    ptr = C.c_void_p()
    rc = lib().sphb_host_alloc(C.byref(ptr), n * PARTICLE.itemsize)
    if rc != 0:
        raise RuntimeError(f"sphb_host_alloc failed: {rc}")
    buf = (C.c_char * (n * PARTICLE.itemsize)).from_address(ptr.value)
    return np.frombuffer(buf, dtype=PARTICLE), ptr


def jitter_lattice(stage, case, amount):
    """SURVEY.md 8(d) config 4's non-lattice variant: every particle displaced by uniform
    +-amount lattice spacings per axis, mt19937(seed = 12345)."""
    n = case["n"]
    spacing = float(np.float32(np.cbrt(np.float32(case["volume"]))) / np.float32(np.ceil(np.cbrt(np.float32(n)))))
    rng = np.random.Generator(np.random.MT19937(12345))
    for lo in range(0, n, 1 << 22):  # in pieces: 16.7 M x 3 doubles at once would be 400 MB
        hi = min(n, lo + (1 << 22))
        d = (rng.random((hi - lo, 3), dtype=np.float32) * np.float32(2) - np.float32(1)) * np.float32(amount * spacing)
        stage["position"][lo:hi, :3] += d


def run_gpu_case(case, steps, warmup, device, e2e_steps, sampler=None, options=(), jitter=0.0):
    """Times `steps` resident steps (after `warmup`), then `e2e_steps` host-buffer steps."""
    from libclsph_b200 import Engine, PASS_NAMES, host
    n = case["n"]
    sc = case["scene"]
    stage, stage_ptr = pinned_particles(n)
    host.lattice(n, case["volume"], out=stage)
    if jitter:
        jitter_lattice(stage, case, jitter)
    e = Engine(case["params"], case["terms"], device)
    for k, v in options:
        e.set_option(k, v)
    e.set_scene(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
    e.upload_ptr(stage_ptr.value, n)
    dt = 1.0 / 600.0
    # the timed steps go through sphb_run_frame, the call sph_simulation::simulate() makes for a
    # frame's sub-steps, in its default mode (device-driven CUDA-graph steps up to 6 M particles,
    # one host round trip per step above)
    dt, _, _ = e.run_frame(dt, 1.0, warmup)
    l0 = e.launch_count()
    if sampler:
        sampler.start()
    e.event_record(0)
    dt, _, done = e.run_frame(dt, 1.0, steps)
    e.event_record(1)
    assert done == steps
    ms = e.event_elapsed_ms(0, 1)
    clocks = sampler.finish() if sampler else None
    launches = e.launch_count() - l0
    # per-pass CUDA-event times come from `steps` FURTHER steps, so that the timed region above
    # carries no profiling events; the reference-rank kernels, which otherwise run on a side
    # stream beside the density pass, are kept on the main stream there, so that every pass
    # time is that pass's own kernels' time
    e.set_option("side_stream", 0)
    e.profile(True)
    e.event_record(0)
    for _ in range(steps):
        dt = e.step(dt)
    e.event_record(1)
    ms_serial = e.event_elapsed_ms(0, 1)
    pass_ms, pass_launches, psteps = e.pass_times()
    e.profile(False)
    e.set_option("side_stream", 1)
    params_now = e.parameters()
    list_stats = e.neighbor_list_stats() if not any(k == "neighbor_kernel" and v for k, v in options) else None
    res = dict(ms_total=ms, ms_profiled_serial=ms_serial, launches=launches, pass_ms=pass_ms, pass_launches=pass_launches,
               list_stats=list_stats,
               clocks=clocks, dt=dt, grid=[int(params_now[k][0]) for k in ("grid_size_x", "grid_size_y", "grid_size_z")],
               grid_cell_count=int(params_now["grid_cell_count"][0]))
    # the frame loop as one engine call (sphb_run_frame): the same steps enqueued without a host
    # round trip per step, against the per-step path, `steps` steps each
    res["frame_loop_ms_per_step"] = {}
    for name, opts in (("host_round_trip_per_step", (("device_loop", 0),)),
                       ("device_loop", (("device_loop", 1), ("graph", 0))),
                       ("device_loop_graph", (("device_loop", 1), ("graph", 1)))):
        for k, v in opts:
            e.set_option(k, v)
        # every mode runs the same trajectory: the lattice again, `warmup` steps, `steps` timed steps
        e.upload_ptr(stage_ptr.value, n)
        dt, _, _ = e.run_frame(1.0 / 600.0, 1.0, warmup)
        e.event_record(0)
        dt, _, done = e.run_frame(dt, 1.0, steps)
        e.event_record(1)
        res["frame_loop_ms_per_step"][name] = e.event_elapsed_ms(0, 1) / max(1, done)
    res["queued_steps"] = e.queued_steps()
    e.set_option("graph", 1)
    e.set_option("device_loop", 2)
    # end to end: host buffers in, host buffers out, every step
    if e2e_steps > 0:
        e.download_ptr(stage_ptr.value)
        for _ in range(1):  # one warm pass of the host path
            e.upload_ptr(stage_ptr.value, n)
            dt = e.step(dt)
            e.download_ptr(stage_ptr.value)
        t0 = time.perf_counter()
        for _ in range(e2e_steps):
            e.upload_ptr(stage_ptr.value, n)
            dt = e.step(dt)
            e.download_ptr(stage_ptr.value)
        res["e2e_serial_s"] = time.perf_counter() - t0
        res["e2e_steps"] = e2e_steps
        # the same three things per step — 80 n bytes up from pinned memory, one step, 80 n bytes
        # back — over INDEPENDENT batches (every step starts from the host's lattice), so that the
        # upload of batch k+1 and the read-back of batch k-1 run beside step k (sphb_upload_begin /
        # _commit, sphb_download_begin / _wait; two pinned host arrays)
        src, src_ptr = pinned_particles(n)
        host.lattice(n, case["volume"], out=src)
        dt0 = 1.0 / 600.0
        for timed in (False, True):
            k_steps = e2e_steps if timed else 1
            e.download_wait()
            if timed:
                t0 = time.perf_counter()
            e.upload_begin(src_ptr.value, n)
            for k in range(k_steps):
                e.upload_commit()
                if k + 1 < k_steps:
                    e.upload_begin(src_ptr.value, n)
                e.step(dt0)
                e.download_wait()
                e.download_begin(stage_ptr.value)
            e.download_wait()
            if timed:
                res["e2e_s"] = time.perf_counter() - t0
        from libclsph_b200 import lib as _lib
        _lib().sphb_host_free(src_ptr)
        del src
        # the pattern sph_simulation::simulate() itself has (sph_simulation.cpp:374-427): the state
        # stays on the device, every frame of 10 sub-steps ends with ONE read-back of all particles
        # into the host array (here overlapped with the next frame's steps, as simulate() does when
        # only save_frame is hooked)
        frames, substeps = max(2, e2e_steps), 10
        e.download_begin(stage_ptr.value)
        t0 = time.perf_counter()
        for _ in range(frames):
            dt, _, _ = e.run_frame(dt, 1.0, substeps)
            e.download_wait()
            e.download_begin(stage_ptr.value)
        e.download_wait()
        res["frames_s"] = time.perf_counter() - t0
        res["frames"] = (frames, substeps)
        res["finite"] = bool(np.isfinite(stage["position"][:: max(1, n // 4096), :3]).all())
    e.close()
    del stage
    import ctypes as C
    from libclsph_b200 import lib
    lib().sphb_host_free(stage_ptr)
    return res


def list_summary(st, n):
    """The neighbour lists of a step, per particle (sphb_neighbor_list_stats)."""
    if not st:
        return None
    pairs = (n + 1) // 2
    return dict(list_overflows=st["overflowed_pairs"], overflow_fraction=st["overflowed_pairs"] / pairs,
                split_pairs_fraction=st["split_pairs"] / pairs, entries_per_pair=st["entries"] / pairs,
                longest_list=st["longest"])


def run_dam_break_horizon(case, device, warmup, horizon=200, window=20):
    """the dam break over steps 0-`horizon`, the first and the last
    `window` steps timed separately (the rest lattice, then the collapsing, splashing block).  The
    engine is warmed up on the same state first, then the lattice is uploaded again."""
    from libclsph_b200 import Engine, host
    n, sc = case["n"], case["scene"]
    stage, stage_ptr = pinned_particles(n)
    host.lattice(n, case["volume"], out=stage)
    e = Engine(case["params"], case["terms"], device)
    e.set_scene(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
    e.upload_ptr(stage_ptr.value, n)
    e.run_frame(1.0 / 600.0, 1.0, warmup)
    e.upload_ptr(stage_ptr.value, n)
    out, dt, done_total = {}, 1.0 / 600.0, 0
    for name, k, timed in ((f"steps_0_{window}", window, True), ("between", horizon - 2 * window, False),
                           (f"steps_{horizon - window}_{horizon}", window, True)):
        e.event_record(0)
        dt, _, done = e.run_frame(dt, 1.0, k)
        e.event_record(1)
        done_total += done
        if timed:
            ms = e.event_elapsed_ms(0, 1)
            out[name] = dict(ms_per_step=ms / max(1, done), value=n * done / (ms * 1e-3), steps=done,
                             neighbor_lists=list_summary(e.neighbor_list_stats(), n), dt=dt)
    p = e.parameters()
    out["grid_after"] = [int(p[k][0]) for k in ("grid_size_x", "grid_size_y", "grid_size_z")]
    out["steps_done"] = done_total
    e.download_ptr(stage_ptr.value)
    out["finite"] = bool(np.isfinite(stage["position"][:: max(1, n // 4096), :3]).all())
    e.close()
    del stage
    from libclsph_b200 import lib
    lib().sphb_host_free(stage_ptr)
    return out


def cpu_arm(budget_s, steps, warmup, kind_pref="reference", fixed_n=None):
    """The reference's CPU implementation of the step (oracle/_ref = its own kernels/sph.cl +
    scene.cpp compiled as C++, OpenMP over work-items; else the plain-C port) on EVERY host core
    this process may run on (its affinity mask — not OMP_NUM_THREADS, which torch.distributed.run
    sets to 1).  fixed_n: time exactly that many particles (the headline configuration); else a
    bounded water-box sample sized so that (steps + warmup) steps take about budget_s seconds."""
    from oracle import pyorc
    kind = kind_pref if pyorc.available(kind_pref) else "port"
    if kind == "port" and not pyorc.available("port"):
        pyorc.build("port")
    orc = pyorc.load(kind)
    orc.lib.orc_set_threads(orc.cores)
    cores = orc.lib.orc_get_threads()

    def prepare(n, tmp):
        p, t, vol = orc.settings(WATER, dict(SIM, particles_count=int(n)))
        dims = water_box_dims(n)
        os.makedirs(os.path.join(tmp, "scenes"), exist_ok=True)
        write_box_obj(os.path.join(tmp, "scenes", f"box_{n}.obj"), *dims)
        if kind == "reference":
            sc = orc.ref_scene(tmp, f"box_{n}.obj", float(p["h"][0]) * 2)
        else:
            from libclsph_b200 import host
            sc = host.load_scene(os.path.join(tmp, "scenes", f"box_{n}.obj"), float(p["h"][0]) * 2)
        df = orc.compute_df(sc["bbs"], sc["transforms"], sc["rvertices"], int(sc["total_gridpoints"]))
        return p, t, orc.init_particles(p, vol), sc, df

    def run(n, k, w, tmp):
        p, t, front, sc, df = prepare(n, tmp)
        dt = 1.0 / 600.0
        for _ in range(w):
            dt = orc.step(front, p, t, df, sc["bbs"], dt)["dt"]
        t0 = time.perf_counter()
        for _ in range(k):
            dt = orc.step(front, p, t, df, sc["bbs"], dt)["dt"]
        return time.perf_counter() - t0

    with tempfile.TemporaryDirectory() as tmp:
        if fixed_n:
            n, side = int(fixed_n), None
        else:
            probe_n = 32 ** 3
            rate = probe_n / max(run(probe_n, 1, 1, tmp), 1e-6)  # particle-steps/s, calibration
            want = rate * budget_s / max(1, steps + warmup)
            side = int(max(32, min(160, np.floor(np.cbrt(want)))))
            n = side ** 3
        secs = run(n, steps, warmup, tmp)
    shape = f" ({side}^3)" if side else ""
    return dict(value=n * steps / secs, unit="particle-steps/s", cores=int(cores), kind=kind,
                sample=f"{n} particle{shape} water box, {steps} step(s) after {warmup} warm-up, "
                       f"{secs:.1f} s on {cores} host threads; same physics and scene family as the GPU workload",
                n=n, seconds=secs, steps=steps, warmup=warmup)


def emit(line):
    """The ONE JSON line, on the real stdout (fd 1 is pointed at stderr while the run is going,
    because native code on the path prints progress lines)."""
    os.write(_REAL_STDOUT, (json.dumps(line) + "\n").encode())


_REAL_STDOUT = 1


def main():
    global _REAL_STDOUT
    sys.stdout.flush()
    _REAL_STDOUT = os.dup(1)
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--particles", type=int, default=16777216)
    ap.add_argument("--e2e-steps", type=int, default=3)
    ap.add_argument("--cpu-budget", type=float, default=20.0, help="seconds of CPU baseline work")
    ap.add_argument("--scaling", default="weak", choices=["weak", "strong"],
                    help="N > 1: one --particles block per GPU (weak) or the N = 1 workload cut in N (strong)")
    ap.add_argument("--scene", default="box", choices=["box", "labyrinth", "river", "shower", "monkey"],
                    help="N > 1, --scaling strong: a reference scene instead of the open-top box")
    ap.add_argument("--driver", default="native", choices=["native", "python"],
                    help="N > 1: the library's own NCCL step driver, or the Python protocol driver")
    ap.add_argument("--p2p", type=int, default=1, help="N > 1, native driver: 1 = halo exchange over peer "
                    "memory when it can be mapped, 0 = ncclSend / ncclRecv")
    ap.add_argument("--no-parity", action="store_true", help="N > 1: skip the parity check that precedes the timing")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg")
    ap.add_argument("--no-secondary", action="store_true", help="skip the 1 M dam-break leg")
    args = ap.parse_args()
    rank = int(os.environ.get("RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    warmup = max(args.warmup, 3) if args.impl == "ours" else args.warmup
    workload = f"synthetic {args.particles}-particle water box (water.json + default.json physics, " \
               f"reference lattice, open-top box), BASELINE.json configs[3]"

    if args.impl == "reference":
        if rank != 0:
            return 0
        # The headline configuration itself: ONE timed step at --particles after ONE warm-up step
        # (16.7 M: ~25 s per step on 16 cores), whatever --steps / --warmup ask for — a CPU step is
        # the same work every time, and K + W of them would take the better part of an hour.  At
        # N > 1 the GPU arm's workload is N such blocks; the CPU arm still times one block (its
        # particle-steps/s do not depend on how many blocks stand side by side) and says so.
        full = cpu_arm(0.0, 1, 1, fixed_n=args.particles)
        small = cpu_arm(0.0, 2, 1, fixed_n=160 ** 3) if args.particles > 160 ** 3 else None
        same = args.gpus == 1
        line = dict(metric="SPH particle-steps/sec", value=full["value"], unit="particle-steps/s", impl="reference",
                    n_gpus=args.gpus, steps=full["steps"], warmup=full["warmup"],
                    requested=dict(steps=args.steps, warmup=args.warmup),
                    ms_per_step=1e3 * full["seconds"] / full["steps"], higher_is_better=True, scaling="weak",
                    vs_baseline=None, dtype="f32", data="synthetic",
                    config=dict(workload=workload, timed_sample=full["sample"], same_config=same,
                                same_config_note=None if same else
                                f"the GPU arm steps {args.gpus} such blocks side by side; the CPU arm times one",
                                second_sample=small and dict(value=small["value"], sample=small["sample"])),
                    cpu_baseline=dict(value=full["value"], unit=full["unit"], cores=full["cores"], kind=full["kind"],
                                      sample=full["sample"]),
                    opencl_cpu_runtime=opencl_cpu_runtime(),
                    e2e=dict(value=full["value"], unit="particle-steps/s", h2d_bytes_per_step=0, d2h_bytes_per_step=0))
        emit(line)
        return 0

    if world > 1 or args.gpus > 1:
        from bench_multi import main_multi
        return main_multi(args, rank, world, local, workload)

    from libclsph_b200 import build
    build.build_all()
    with tempfile.TemporaryDirectory() as tmp:
        case = build_case(tmp, args.particles, "water_box")
        sampler = ClockSampler(physical_device(local))
        r = run_gpu_case(case, args.steps, warmup, local, args.e2e_steps, sampler)
        secondary = None
        if not args.no_secondary:
            c2 = build_case(tmp, 1000000, "dam_break")
            r2 = run_gpu_case(c2, max(args.steps, 20), warmup, local, 0)
            secondary = dict(workload="synthetic 1000000-particle water dam break, BASELINE.json configs[2]",
                             value=c2["n"] * max(args.steps, 20) / (r2["ms_total"] * 1e-3), unit="particle-steps/s",
                             ms_per_step=r2["ms_total"] / max(args.steps, 20),
                             pass_ms_per_step={k: v / max(args.steps, 20) for k, v in r2["pass_ms"].items()},
                             frame_loop_ms_per_step=r2["frame_loop_ms_per_step"],
                             neighbor_lists=list_summary(r2["list_stats"], c2["n"]),
                             horizon_0_200=run_dam_break_horizon(c2, local, warmup))
            # SURVEY 8(d) config 4's non-lattice variant: the headline box with every particle
            # displaced by +-0.1 spacings (mt19937(12345)) — irregular neighbour counts
            rj = run_gpu_case(case, args.steps, warmup, local, 0, jitter=0.1)
            secondary["jittered_16M"] = dict(
                workload=f"the headline {case['n']}-particle box, every particle displaced by uniform +-0.1 lattice "
                         "spacings per axis (mt19937 seed 12345), SURVEY.md 8(d) config 4",
                value=case["n"] * args.steps / (rj["ms_total"] * 1e-3), unit="particle-steps/s",
                ms_per_step=rj["ms_total"] / args.steps,
                pass_ms_per_step={k: v / args.steps for k, v in rj["pass_ms"].items() if v},
                neighbor_lists=list_summary(rj["list_stats"], case["n"]))
            # collision-heavy leg (SURVEY 8d config 5's size, config 2's fluid and mesh): the
            # advection pass with the triangle bin, and with the reference's all-faces scan
            c3 = build_case(tmp, 4096000, "mucus_monkeybox")
            k3 = max(args.steps, 10)
            legs = {}
            for name, opt in (("triangle_bin", 1), ("all_faces_scan", 0)):
                r3 = run_gpu_case(c3, k3, warmup, local, 0, options=(("triangle_bin", opt),))
                legs[name] = dict(ms_per_step=r3["ms_total"] / k3, value=c3["n"] * k3 / (r3["ms_total"] * 1e-3),
                                  advect_ms_per_launch=r3["pass_ms"]["advect"] / max(1, r3["pass_launches"]["advect"]) * 2,
                                  pass_ms_per_step={k: v / k3 for k, v in r3["pass_ms"].items() if v},
                                  neighbor_lists=list_summary(r3["list_stats"], c3["n"]))
            secondary["collision_heavy"] = dict(
                workload="synthetic 4096000-particle mucus.json lattice in the 978-face monkeybox scene "
                         "(BASELINE.json configs[1] physics and mesh at configs[4]'s particle count)",
                unit="particle-steps/s", **legs)
    n = case["n"]
    ms_step = r["ms_total"] / args.steps
    value = n * args.steps / (r["ms_total"] * 1e-3)
    peak, peak_src = measured_peak()
    # dominant kernel: whichever of the two neighbour passes takes longer (one launch per step each)
    # (time per STEP: the forces pass is two launches, k_forces_union and the few-microsecond k_forces_overflow)
    dom = max(("density", "forces"), key=lambda k: r["pass_ms"][k])
    f_ms = r["pass_ms"][dom] / args.steps
    achieved = ALGO_BYTES[dom] * n / (f_ms * 1e-3) / 1e9
    # ncu counters of that kernel, taken on THESE sources (else null + why): DRAM traffic per launch,
    # and the two other roofs — issue slots (one warp instruction per clock and SM sub-partition)
    # and the L1 data pipe (one wavefront per clock and SM) — at the clock this run measured
    counters, stale = ncu_counters(dom, n)
    traffic = counters["dram_bytes"] if counters else None
    clock_hz = 1e6 * float((r["clocks"] or {}).get("sm_mhz") or 0.0)
    sms = 148
    try:
        import pynvml
        pynvml.nvmlInit()
        hdl = pynvml.nvmlDeviceGetHandleByIndex(physical_device(local))
        sms = int(pynvml.nvmlDeviceGetNumGpuCores(hdl)) // 128 or 148
    except Exception:
        pass
    other_roofs = {}
    if counters and clock_hz > 0:
        t = f_ms * 1e-3
        other_roofs = dict(
            issue_slots=dict(achieved=counters["warp_inst"] / t, peak=sms * 4 * clock_hz, unit="warp instructions/s",
                             frac=counters["warp_inst"] / t / (sms * 4 * clock_hz),
                             lane_efficiency=counters.get("lane_efficiency"),
                             thread_instructions_per_particle=counters.get("thread_inst", 0.0) / n),
            l1_data_pipe=dict(achieved=counters["l1_wavefronts"] / t, peak=sms * clock_hz, unit="wavefronts/s",
                              frac=counters["l1_wavefronts"] / t / (sms * clock_hz)),
            fp32=dict(note="thread instructions x 2 flop as an upper bound of the fp32 work", unit="TFLOP/s",
                      achieved=2 * counters.get("thread_inst", 0.0) / t / 1e12, peak=sms * 128 * 2 * clock_hz / 1e12,
                      frac=counters.get("thread_inst", 0.0) / t / (sms * 128 * clock_hz)))
    binding = max(other_roofs, key=lambda k: other_roofs[k]["frac"]) if other_roofs else None
    if binding and other_roofs[binding]["frac"] < achieved / peak:
        binding = "hbm"
    radix_passes = (max(1, r["grid_cell_count"].bit_length()) + 7) // 8
    step_bytes = STEP_BYTES_FUSED.get(radix_passes, 301)
    per_pass = {}
    for k, v in r["pass_ms"].items():
        per_pass[k] = dict(ms_per_step=v / args.steps, launches_per_step=r["pass_launches"][k] / args.steps)
        ab = ALGO_BYTES.get(k) if k != "sort" else 16 * radix_passes + 4
        if ab and v > 0:
            per_pass[k]["algo_gbs"] = ab * n / (v / args.steps * 1e-3) / 1e9
    # ncu's DRAM traffic of every captured kernel of a pass (same staleness rule as `traffic`)
    pass_kernels = dict(sort=("k_locate_count", "k_place", "k_order_cells"), reorder=("k_reorder",),
                        density=("k_density_union",), forces=("k_forces_union",), advect=("k_advect",),
                        refrank=("k_morton_hist", "k_reference_rank"))
    try:
        doc = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        if doc.get("source_hash") == kernel_source_hash():
            for k, names in pass_kernels.items():
                recs = {nm: doc["kernels"][nm] for nm in names if nm in doc.get("kernels", {}) and
                        doc["kernels"][nm].get("particles") == n}
                if recs and k in per_pass:
                    per_pass[k]["ncu_dram"] = {nm: dict(gb_per_launch=r_["dram_bytes"] / 1e9, gbs=r_["dram_gbs"],
                                                        frac_of_peak=r_["dram_gbs"] / peak) for nm, r_ in recs.items()}
    except Exception:
        pass
    line = dict(
        metric="SPH particle-steps/sec", value=value, unit="particle-steps/s", n_gpus=1, steps=args.steps,
        warmup=warmup, ms_per_step=ms_step, higher_is_better=True, scaling="weak", vs_baseline=None,
        dtype="f32", data="synthetic",
        config=dict(workload=workload, particles=n, box=[round(x, 4) for x in case["dims"]], grid=r["grid"],
                    radix_passes=radix_passes, l2="inputs exceed L2 (positions alone are %d MB)" % (16 * n >> 20),
                    parallelism="1 GPU",
                    pass_times="CUDA events over %d further steps with the reference-rank kernels kept on the "
                               "main stream (%.3f ms per step that way; the timed steps overlap them with the "
                               "density pass)" % (args.steps, r["ms_profiled_serial"] / args.steps),
                    secondary=secondary),
        clocks=r["clocks"], gpu_launches=r["launches"],
        frame_loop=dict(ms_per_step=r["frame_loop_ms_per_step"], steps_without_host_round_trip=r["queued_steps"],
                        note="sphb_run_frame over %d steps per mode, each from the same state; `value` is "
                             "its default mode (device_loop_graph up to 6 M particles)" % args.steps),
        roofline=dict(bound="hbm", kernel={"density": "density + pressure pass (k_density_union)",
                                           "forces": "forces pass (k_forces_union)"}[dom],
                      achieved=achieved, peak=peak, unit="GB/s", frac=achieved / peak, traffic=traffic,
                      traffic_note=stale or "ncu dram__bytes_read.sum + dram__bytes_write.sum per launch, captured on "
                                            "these kernel sources (profiles/ncu_traffic.json)",
                      peak_source=peak_src, algorithmic_bytes_per_particle=ALGO_BYTES[dom], launch_ms=f_ms,
                      binding_bound=binding, other_roofs=other_roofs,
                      both_passes={k: dict(launch_ms=r["pass_ms"][k] / args.steps,
                                           frac=ALGO_BYTES[k] * n / (r["pass_ms"][k] / args.steps * 1e-3) / 1e9 / peak)
                                   for k in ("density", "forces")},
                      note="`frac` is the HBM fraction the contract asks for (algorithmic bytes / time / measured "
                           "copy peak); what BINDS the kernel is `binding_bound` (DESIGN.md section 3): the passes "
                           "are bound by the L1 data pipe, not by HBM",
                      whole_step=dict(algorithmic_bytes_per_particle=step_bytes,
                                      achieved=step_bytes * n / (ms_step * 1e-3) / 1e9,
                                      frac=step_bytes * n / (ms_step * 1e-3) / 1e9 / peak)),
        passes=per_pass, neighbor_lists=list_summary(r["list_stats"], n), opencl_cpu_runtime=opencl_cpu_runtime())
    if "e2e_s" in r:
        line["e2e"] = dict(value=n * r["e2e_steps"] / r["e2e_serial_s"], unit="particle-steps/s",
                           h2d_bytes_per_step=80 * n, d2h_bytes_per_step=80 * n + 32, steps=r["e2e_steps"],
                           finite=r["finite"],
                           note="a chain of DEPENDENT steps through host memory, through the C ABI: every step uploads "
                                "all particles (80-byte AoS) from pinned host memory, steps once, and reads all "
                                "particles back, each waiting for the one before",
                           pipelined=dict(value=n * r["e2e_steps"] / r["e2e_s"], unit="particle-steps/s",
                                          note="the same three things per step over INDEPENDENT batches, so that the "
                                               "upload of the next batch and the read-back of the previous one "
                                               "overlap the step (sphb_upload_begin/_commit, "
                                               "sphb_download_begin/_wait)"),
                           simulate_frames=dict(
                               value=n * r["frames"][0] * r["frames"][1] / r["frames_s"], unit="particle-steps/s",
                               note="the copy pattern of sph_simulation::simulate(): state resident, %d frames of %d "
                                    "sub-steps (sphb_run_frame), one %d-byte read-back per frame overlapped with the "
                                    "next frame's steps" % (r["frames"][0], r["frames"][1], 80 * n),
                               d2h_bytes_per_step=80 * n // r["frames"][1], h2d_bytes_per_step=0))
    if not args.no_cpu:
        cb = cpu_arm(args.cpu_budget, 2, 1)
        line["cpu_baseline"] = dict(value=cb["value"], unit=cb["unit"], cores=cb["cores"], kind=cb["kind"],
                                    sample=cb["sample"])
    emit(line)
    return 0


if __name__ == "__main__":
    sys.exit(main())
