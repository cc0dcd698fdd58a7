"""ctypes binding of include/sphb.h (one engine = one GPU = one stream)."""
import ctypes as C
import os

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

PARTICLE = np.dtype(
    [
        ("position", np.float32, 4),
        ("velocity", np.float32, 4),
        ("intermediate_velocity", np.float32, 4),
        ("acceleration", np.float32, 4),
        ("density", np.float32),
        ("pressure", np.float32),
        ("grid_index", np.uint32),
        ("pad_", np.uint32),
    ]
)
PARAMS = np.dtype(
    [
        ("particles_count", np.uint32),
        ("fluid_density", np.float32),
        ("total_mass", np.float32),
        ("particle_mass", np.float32),
        ("dynamic_viscosity", np.float32),
        ("simulation_time", np.float32),
        ("h", np.float32),
        ("simulation_scale", np.float32),
        ("target_fps", np.float32),
        ("surface_tension_threshold", np.float32),
        ("surface_tension", np.float32),
        ("restitution", np.float32),
        ("K", np.float32),
        ("pad0_", np.uint32, 3),
        ("constant_acceleration", np.float32, 4),
        ("grid_size_x", np.int32),
        ("grid_size_y", np.int32),
        ("grid_size_z", np.int32),
        ("grid_cell_count", np.uint32),
        ("min_point", np.float32, 4),
        ("max_point", np.float32, 4),
    ]
)
TERMS = np.dtype(
    [
        ("poly_6", np.float32),
        ("poly_6_gradient", np.float32),
        ("poly_6_laplacian", np.float32),
        ("spiky", np.float32),
        ("viscosity", np.float32),
    ]
)
BB = np.dtype(
    [
        ("maxx", np.float32),
        ("maxy", np.float32),
        ("maxz", np.float32),
        ("minx", np.float32),
        ("miny", np.float32),
        ("minz", np.float32),
        ("size_x", np.uint64),
        ("size_y", np.uint64),
        ("size_z", np.uint64),
        ("offset", np.uint64),
    ]
)
assert (PARTICLE.itemsize, PARAMS.itemsize, TERMS.itemsize, BB.itemsize) == (80, 128, 20, 56)

PASS_NAMES = ["bbox", "locate", "sort", "reorder", "celltable", "density", "forces", "advect", "refrank"]

# every symbol include/sphb.h declares (tests check the library exports each one)
SYMBOLS = [
    "sphb_create", "sphb_destroy", "sphb_set_scene", "sphb_get_distance_field", "sphb_upload",
    "sphb_step", "sphb_run_frame", "sphb_queued_steps", "sphb_get_parameters", "sphb_download", "sphb_download_begin", "sphb_download_wait",
    "sphb_download_debug",
    "sphb_profile_enable", "sphb_profile_reset", "sphb_get_pass_times", "sphb_launch_count",
    "sphb_stream", "sphb_set_option", "sphb_last_error", "sphb_error_string",
    "sphb_host_alloc", "sphb_host_free", "sphb_debug_advect", "sphb_event_record", "sphb_event_elapsed_ms",
    "sphb_slab_create", "sphb_slab_message_bytes", "sphb_slab_configure", "sphb_slab_upload", "sphb_slab_bbox",
    "sphb_slab_pack", "sphb_slab_unpack", "sphb_slab_forces", "sphb_slab_advect", "sphb_slab_commit",
    "sphb_slab_download", "sphb_compute_dt", "sphb_scene_build_ms", "sphb_upload_begin", "sphb_upload_commit",
    "sphb_nccl_unique_id", "sphb_slab_comm_init", "sphb_slab_start", "sphb_slab_step", "sphb_slab_bytes_sent",
    "sphb_slab_phase_ms", "sphb_slab_resends", "sphb_slab_p2p_active", "sphb_cell_table_noalloc", "sphb_neighbor_list_stats", "sphb_slab_gpu_segments",
]


class SphbError(RuntimeError):
    def __init__(self, code, text):
        super().__init__(f"[{code}] {text}")
        self.code = code


def lib_path():
    # SPHB_LIB: an engine built with other compile-time parameters (experiments, tools/ab_frame.py)
    return os.environ.get("SPHB_LIB") or os.path.join(HERE, "libsphb.so")


_lib = None


def lib():
    """Loads libsphb.so.  There is no fallback: a missing library is an error."""
    global _lib
    if _lib is None:
        path = lib_path()
        if not os.path.exists(path):
            raise ImportError(
                f"{path} is missing: build it with `python -m libclsph_b200.build` "
                "(the sphb engine has no CPU or pure-Python path)")
        L = C.CDLL(path)
        vp, u32, i32, f32 = C.c_void_p, C.c_uint32, C.c_int, C.c_float
        L.sphb_create.argtypes = [vp, vp, i32, C.POINTER(vp)]
        L.sphb_destroy.argtypes = [vp]
        L.sphb_set_scene.argtypes = [vp, vp, u32, vp, vp, u32]
        L.sphb_get_distance_field.argtypes = [vp, vp]
        L.sphb_upload.argtypes = [vp, vp, u32]
        L.sphb_upload_begin.argtypes = [vp, vp, u32]
        L.sphb_upload_commit.argtypes = [vp]
        L.sphb_step.argtypes = [vp, f32, C.POINTER(f32)]
        L.sphb_run_frame.argtypes = [vp, f32, f32, C.c_uint64, C.POINTER(f32), C.POINTER(u32), vp, u32]
        L.sphb_queued_steps.argtypes = [vp]
        L.sphb_queued_steps.restype = C.c_uint64
        L.sphb_get_parameters.argtypes = [vp, vp]
        L.sphb_download.argtypes = [vp, vp]
        L.sphb_download_begin.argtypes = [vp, vp]
        L.sphb_download_wait.argtypes = [vp]
        L.sphb_download_debug.argtypes = [vp, vp, vp, vp, u32, C.POINTER(i32)]
        L.sphb_profile_enable.argtypes = [vp, i32]
        L.sphb_profile_reset.argtypes = [vp]
        L.sphb_get_pass_times.argtypes = [vp, vp, vp, C.POINTER(u32)]
        L.sphb_launch_count.argtypes = [vp]
        L.sphb_launch_count.restype = C.c_uint64
        L.sphb_stream.argtypes = [vp]
        L.sphb_stream.restype = vp
        L.sphb_set_option.argtypes = [vp, C.c_char_p, i32]
        L.sphb_last_error.argtypes = [vp]
        L.sphb_last_error.restype = C.c_char_p
        L.sphb_error_string.argtypes = [i32]
        L.sphb_error_string.restype = C.c_char_p
        L.sphb_slab_create.argtypes = [vp, vp, i32, u32, u32, C.POINTER(vp)]
        L.sphb_slab_message_bytes.argtypes = [vp]
        L.sphb_slab_message_bytes.restype = C.c_size_t
        L.sphb_slab_configure.argtypes = [vp, f32, f32, i32, i32]
        L.sphb_slab_upload.argtypes = [vp, vp, vp, u32]
        L.sphb_slab_bbox.argtypes = [vp, vp]
        L.sphb_slab_pack.argtypes = [vp, vp, vp]
        L.sphb_slab_unpack.argtypes = [vp, vp, vp, C.POINTER(u32), C.POINTER(u32)]
        L.sphb_slab_forces.argtypes = [vp, vp, vp]
        L.sphb_slab_advect.argtypes = [vp, f32, vp]
        L.sphb_slab_commit.argtypes = [vp]
        L.sphb_slab_download.argtypes = [vp, vp, vp, u32, C.POINTER(u32)]
        L.sphb_nccl_unique_id.argtypes = [vp]
        L.sphb_slab_comm_init.argtypes = [vp, vp, i32, i32]
        L.sphb_slab_start.argtypes = [vp]
        L.sphb_slab_step.argtypes = [vp, f32, C.POINTER(f32), C.POINTER(i32), C.POINTER(u32), C.POINTER(u32)]
        L.sphb_slab_bytes_sent.argtypes = [vp]
        L.sphb_slab_bytes_sent.restype = C.c_uint64
        L.sphb_slab_phase_ms.argtypes = [vp, vp, i32]
        L.sphb_slab_gpu_segments.argtypes = [vp, vp, vp, i32]
        L.sphb_slab_p2p_active.argtypes = [vp]
        L.sphb_slab_resends.argtypes = [vp]
        L.sphb_slab_resends.restype = C.c_uint64
        L.sphb_compute_dt.argtypes = [vp, f32, f32]
        L.sphb_compute_dt.restype = f32
        L.sphb_scene_build_ms.argtypes = [vp]
        L.sphb_scene_build_ms.restype = f32
        L.sphb_debug_advect.argtypes = [vp, f32]
        L.sphb_event_record.argtypes = [vp, i32]
        L.sphb_event_elapsed_ms.argtypes = [vp, i32, i32, C.POINTER(f32)]
        L.sphb_host_alloc.argtypes = [C.POINTER(vp), C.c_size_t]
        L.sphb_host_free.argtypes = [vp]
        L.sphb_cell_table_noalloc.argtypes = [vp, u32, vp, vp, u32, vp]
        L.sphb_neighbor_list_stats.argtypes = [vp, vp]
        _lib = L
    return _lib


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


class Engine:
    """One sphb_engine.  Arrays are numpy structured arrays of the dtypes above."""

    def __init__(self, params, terms, device=-1, slab_capacity=None, exchange_capacity=0):
        """slab_capacity=None: a whole-domain engine (sphb_create); else a slab engine
        (sphb_slab_create) with that many local particle slots."""
        self.L = lib()
        self.h = C.c_void_p()
        params = np.ascontiguousarray(params, PARAMS)
        terms = np.ascontiguousarray(terms, TERMS)
        if slab_capacity is None:
            rc = self.L.sphb_create(_ptr(params), _ptr(terms), device, C.byref(self.h))
        else:
            rc = self.L.sphb_slab_create(_ptr(params), _ptr(terms), device, int(slab_capacity),
                                         int(exchange_capacity), C.byref(self.h))
        self.params0 = params.copy()
        if rc != 0:
            self.h = None
            raise SphbError(rc, self.L.sphb_error_string(rc).decode())
        self.n = int(params.reshape(-1)[0]["particles_count"])

    def _check(self, rc):
        if rc != 0:
            raise SphbError(rc, self.L.sphb_last_error(self.h).decode())

    def close(self):
        if getattr(self, "h", None):
            self.L.sphb_destroy(self.h)
            self.h = None

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def set_scene(self, bbs, transforms, rvertices, total_grid_points):
        bbs = np.ascontiguousarray(bbs, BB)
        transforms = np.ascontiguousarray(transforms, np.float32)
        rvertices = np.ascontiguousarray(rvertices, np.float32)
        self.total_grid_points = int(total_grid_points)
        self._check(self.L.sphb_set_scene(self.h, _ptr(bbs), len(bbs), _ptr(transforms), _ptr(rvertices),
                                          self.total_grid_points))

    def scene_build_ms(self):
        """device milliseconds of the last distance-field build"""
        return float(self.L.sphb_scene_build_ms(self.h))

    def distance_field(self):
        out = np.zeros(self.total_grid_points, np.float32)
        self._check(self.L.sphb_get_distance_field(self.h, _ptr(out)))
        return out

    def upload(self, particles):
        particles = np.ascontiguousarray(particles, PARTICLE)
        self._check(self.L.sphb_upload(self.h, _ptr(particles), len(particles)))

    def upload_begin(self, ptr, n):
        """starts the host -> device copy of n particles at `ptr` (pinned memory); see upload_commit"""
        self._check(self.L.sphb_upload_begin(self.h, C.c_void_p(ptr), n))

    def upload_commit(self):
        self._check(self.L.sphb_upload_commit(self.h))

    def upload_ptr(self, ptr, n):
        """ptr: address of n 80-byte particles in (ideally pinned) host memory."""
        self._check(self.L.sphb_upload(self.h, C.c_void_p(ptr), n))

    def step(self, dt):
        out = C.c_float()
        self._check(self.L.sphb_step(self.h, dt, C.byref(out)))
        return out.value

    def run_frame(self, dt, frame_time, max_steps=0, log_cap=4096):
        """The sub-step loop of one frame on the device -> (dt afterwards, timeleft after each step)."""
        out = C.c_float()
        steps = C.c_uint32()
        log = np.zeros(log_cap, np.float32)
        self._check(self.L.sphb_run_frame(self.h, dt, frame_time, max_steps, C.byref(out), C.byref(steps),
                                          _ptr(log), log_cap))
        return out.value, log[:min(steps.value, log_cap)].copy(), steps.value

    def queued_steps(self):
        return int(self.L.sphb_queued_steps(self.h))

    def parameters(self):
        p = np.zeros(1, PARAMS)
        self._check(self.L.sphb_get_parameters(self.h, _ptr(p)))
        return p

    def download(self, out=None):
        if out is None:
            out = np.zeros(self.n, PARTICLE)
        self._check(self.L.sphb_download(self.h, _ptr(out)))
        return out

    def download_ptr(self, ptr):
        self._check(self.L.sphb_download(self.h, C.c_void_p(ptr)))

    def download_begin(self, ptr):
        """Snapshot the state and start copying it to host address `ptr` (page-locked memory from
        host_alloc for a copy that overlaps later steps); complete after download_wait()."""
        self._check(self.L.sphb_download_begin(self.h, C.c_void_p(ptr)))

    def download_wait(self):
        self._check(self.L.sphb_download_wait(self.h))

    def debug(self):
        """dict(keys, perm, cell_table, attempts) of the last step."""
        ncell = int(self.parameters()["grid_cell_count"][0])
        keys = np.zeros(self.n, np.uint32)
        perm = np.zeros(self.n, np.uint32)
        table = np.zeros(ncell, np.uint32)
        att = C.c_int()
        self._check(self.L.sphb_download_debug(self.h, _ptr(keys), _ptr(perm), _ptr(table), ncell, C.byref(att)))
        return dict(keys=keys, perm=perm, cell_table=table, attempts=att.value)

    def debug_advect(self, dt):
        self._check(self.L.sphb_debug_advect(self.h, dt))

    def profile(self, on=True):
        self._check(self.L.sphb_profile_enable(self.h, int(on)))
        self._check(self.L.sphb_profile_reset(self.h))

    def pass_times(self):
        """(dict pass -> ms accumulated, dict pass -> launches, steps) since profile()."""
        ms = np.zeros(len(PASS_NAMES), np.float32)
        ln = np.zeros(len(PASS_NAMES), np.uint32)
        steps = C.c_uint32()
        self._check(self.L.sphb_get_pass_times(self.h, _ptr(ms), _ptr(ln), C.byref(steps)))
        return dict(zip(PASS_NAMES, ms.tolist())), dict(zip(PASS_NAMES, ln.tolist())), steps.value

    def event_record(self, slot):
        self._check(self.L.sphb_event_record(self.h, slot))

    def event_elapsed_ms(self, a, b):
        ms = C.c_float()
        self._check(self.L.sphb_event_elapsed_ms(self.h, a, b, C.byref(ms)))
        return ms.value

    def neighbor_list_stats(self):
        """dict(overflowed_pairs, split_pairs, entries, longest) of the last step's neighbour lists"""
        out = (C.c_uint64 * 4)()
        self._check(self.L.sphb_neighbor_list_stats(self.h, out))
        return dict(overflowed_pairs=int(out[0]), split_pairs=int(out[1]), entries=int(out[2]), longest=int(out[3]))

    def launch_count(self):
        return int(self.L.sphb_launch_count(self.h))

    def stream(self):
        return self.L.sphb_stream(self.h)

    def set_option(self, name, value):
        self._check(self.L.sphb_set_option(self.h, name.encode(), int(value)))
