"""ctypes binding of include/clsph_host.h: the C++11 host classes (scene, sph_simulation,
houdini_file_saver) behind C entry points.  No fallback: a missing library is an error."""
import ctypes as C
import os

import numpy as np

from .engine import BB, PARAMS, PARTICLE, TERMS, HERE, _ptr, lib as _engine_lib

SYMBOLS = [
    "clsph_scene_load", "clsph_scene_face_count", "clsph_scene_total_grid_points", "clsph_scene_copy",
    "clsph_scene_free", "clsph_load_settings", "clsph_lattice", "clsph_geo_text", "clsph_bgeo_bytes", "clsph_format_floats",
    "clsph_save_frame", "clsph_simulate", "clsph_simulate_devices", "clsph_equal_count_cuts", "clsph_gather_sort",
    "clsph_deal_out",
]

_lib = None


def lib_path():
    # CLSPH_HOST_LIB: another build of the same library (tools/sanitize_cpu.sh: ASan / UBSan)
    return os.environ.get("CLSPH_HOST_LIB") or os.path.join(HERE, "libclsph_host.so")


def lib():
    global _lib
    if _lib is None:
        _engine_lib()  # libsphb.so first (the host library links it)
        path = lib_path()
        if not os.path.exists(path):
            raise ImportError(f"{path} is missing: build it with `python -m libclsph_b200.build`")
        L = C.CDLL(path)
        vp, u32, i32, f32 = C.c_void_p, C.c_uint32, C.c_int, C.c_float
        L.clsph_scene_load.argtypes = [C.c_char_p, f32]
        L.clsph_scene_load.restype = vp
        L.clsph_scene_face_count.argtypes = [vp]
        L.clsph_scene_face_count.restype = u32
        L.clsph_scene_total_grid_points.argtypes = [vp]
        L.clsph_scene_total_grid_points.restype = u32
        L.clsph_scene_copy.argtypes = [vp] * 5
        L.clsph_scene_copy.restype = None
        L.clsph_scene_free.argtypes = [vp]
        L.clsph_scene_free.restype = None
        L.clsph_load_settings.argtypes = [C.c_char_p, C.c_char_p, vp, vp, C.POINTER(f32), C.POINTER(i32),
                                          C.POINTER(i32), C.c_char_p, C.c_size_t]
        L.clsph_lattice.argtypes = [vp, u32, f32]
        L.clsph_lattice.restype = None
        L.clsph_geo_text.argtypes = [vp, u32, f32, u32, vp, C.c_size_t]
        L.clsph_geo_text.restype = C.c_size_t
        L.clsph_bgeo_bytes.argtypes = [vp, vp, u32, vp, C.c_size_t]
        L.clsph_bgeo_bytes.restype = C.c_size_t
        L.clsph_save_frame.argtypes = [C.c_char_p, i32, vp, vp]
        L.clsph_simulate.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, i32, C.c_uint64, C.c_char_p, vp, vp,
                                     C.POINTER(C.c_uint64), C.POINTER(u32), C.c_char_p, C.c_size_t]
        L.clsph_simulate_devices.argtypes = [C.c_char_p, C.c_char_p, C.c_char_p, i32, i32, f32, C.c_uint64,
                                             C.c_char_p, vp, vp, C.POINTER(C.c_uint64), C.POINTER(u32),
                                             C.POINTER(C.c_uint64), C.c_char_p, C.c_size_t]
        _lib = L
    return _lib


def load_scene(obj_path, threshold):
    """scene::load on an OBJ path -> dict(face_count, total_gridpoints, bbs, transforms,
    rvertices, face_normals), or None when the file cannot be loaded."""
    L = lib()
    h = L.clsph_scene_load(os.fsencode(obj_path), float(threshold))
    if not h:
        return None
    try:
        nf = L.clsph_scene_face_count(h)
        sc = dict(face_count=nf, total_gridpoints=L.clsph_scene_total_grid_points(h),
                  bbs=np.zeros(nf, BB), transforms=np.zeros(nf * 12, np.float32),
                  rvertices=np.zeros(nf * 4, np.float32), face_normals=np.zeros(nf * 3, np.float32))
        L.clsph_scene_copy(h, _ptr(sc["bbs"]), _ptr(sc["transforms"]), _ptr(sc["rvertices"]),
                           _ptr(sc["face_normals"]))
        return sc
    finally:
        L.clsph_scene_free(h)


def load_settings(fluid_json, sim_json):
    """sph_simulation::load_settings -> (params, terms, initial_volume, write_all_frames, serialize)"""
    L = lib()
    p = np.zeros(1, PARAMS)
    t = np.zeros(1, TERMS)
    vol = C.c_float()
    w = C.c_int()
    s = C.c_int()
    err = C.create_string_buffer(512)
    rc = L.clsph_load_settings(os.fsencode(fluid_json), os.fsencode(sim_json), _ptr(p), _ptr(t), C.byref(vol),
                               C.byref(w), C.byref(s), err, len(err))
    if rc != 0:
        raise RuntimeError(err.value.decode())
    return p, t, vol.value, bool(w.value), bool(s.value)


def lattice(count, initial_volume, out=None):
    if out is None:
        out = np.zeros(count, PARTICLE)
    lib().clsph_lattice(_ptr(out), count, initial_volume)
    return out


def geo_text(particles, particle_mass, threads=0):
    particles = np.ascontiguousarray(particles, PARTICLE)
    L = lib()
    n = L.clsph_geo_text(_ptr(particles), len(particles), particle_mass, threads, None, 0)
    buf = C.create_string_buffer(n)
    L.clsph_geo_text(_ptr(particles), len(particles), particle_mass, threads, buf, n)
    return buf.raw[:n]


def format_floats(values):
    """the .geo writer's formatter: b"<value>\\n" per float32, as `ostream << float` prints it"""
    values = np.ascontiguousarray(values, np.float32)
    L = lib()
    L.clsph_format_floats.restype = C.c_size_t
    L.clsph_format_floats.argtypes = [C.c_void_p, C.c_size_t, C.c_void_p, C.c_size_t]
    n = L.clsph_format_floats(_ptr(values), len(values), None, 0)
    buf = C.create_string_buffer(n)
    L.clsph_format_floats(_ptr(values), len(values), buf, n)
    return buf.raw[:n]


def bgeo_bytes(particles, params, threads=0):
    """The .bgeo the reference's USE_PARTIO build writes for this frame (params: PARAMS[1])."""
    particles = np.ascontiguousarray(particles, PARTICLE)
    params = np.ascontiguousarray(params, PARAMS)
    assert int(params["particles_count"][0]) == len(particles)
    L = lib()
    n = L.clsph_bgeo_bytes(_ptr(particles), _ptr(params), threads, None, 0)
    buf = C.create_string_buffer(n)
    L.clsph_bgeo_bytes(_ptr(particles), _ptr(params), threads, buf, n)
    return buf.raw[:n]


def save_frame(frames_prefix, particles, params, use_partio=False):
    """houdini_file_saver(prefix).writeFrameToFile(particles, params): frame0000001.geo / .bgeo"""
    particles = np.ascontiguousarray(particles, PARTICLE)
    params = np.ascontiguousarray(params, PARAMS)
    rc = lib().clsph_save_frame(os.fsencode(frames_prefix), int(use_partio), _ptr(particles), _ptr(params))
    if rc != 0:
        raise RuntimeError("clsph_save_frame failed")


def equal_count_cuts(particles, world):
    """the slab faces of sph_simulation::devices (host/sph_simulation_slabs.cpp)"""
    particles = np.ascontiguousarray(particles, PARTICLE)
    out = np.zeros(world + 1, np.float32)
    L = lib()
    L.clsph_equal_count_cuts.argtypes = [C.c_void_p, C.c_uint32, C.c_int, C.c_void_p]
    k = L.clsph_equal_count_cuts(_ptr(particles), len(particles), int(world), _ptr(out))
    if k != world + 1:
        raise RuntimeError("clsph_equal_count_cuts failed")
    return out


def gather_sort(stage, stage_ids, start, cells):
    """the frame gather of sph_simulation::devices (host/sph_simulation_slabs.cpp) on host arrays"""
    stage = np.ascontiguousarray(stage, PARTICLE)
    stage_ids = np.ascontiguousarray(stage_ids, np.uint32)
    start = np.ascontiguousarray(start, np.uint32)
    out, ids = np.zeros(len(stage), PARTICLE), np.zeros(len(stage), np.uint32)
    L = lib()
    L.clsph_gather_sort.argtypes = [C.c_void_p, C.c_void_p, C.c_void_p, C.c_int, C.c_uint32, C.c_void_p, C.c_void_p]
    rc = L.clsph_gather_sort(_ptr(stage), _ptr(stage_ids), _ptr(start), len(start) - 1, int(cells), _ptr(out), _ptr(ids))
    if rc != 0:
        raise RuntimeError("clsph_gather_sort failed")
    return out, ids


def deal_out(particles, ids, cuts, threads):
    """the upload-side partition of sph_simulation::devices (host/sph_simulation_slabs.cpp) on host
    arrays; returns (stage, stage_ids, start)"""
    particles = np.ascontiguousarray(particles, PARTICLE)
    ids = np.ascontiguousarray(ids, np.uint32)
    cuts = np.ascontiguousarray(cuts, np.float32)
    world = len(cuts) - 1
    stage, stage_ids = np.zeros(len(particles), PARTICLE), np.zeros(len(particles), np.uint32)
    start = np.zeros(world + 1, np.uint32)
    L = lib()
    L.clsph_deal_out.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_void_p, C.c_int, C.c_int, C.c_void_p,
                                 C.c_void_p, C.c_void_p]
    rc = L.clsph_deal_out(_ptr(particles), _ptr(ids), len(particles), _ptr(cuts), world, int(threads), _ptr(stage),
                          _ptr(stage_ids), _ptr(start))
    if rc != 0:
        raise RuntimeError("clsph_deal_out failed")
    return stage, stage_ids, start


def simulate(fluid_json, sim_json, obj_path, device=-1, max_steps=0, frames_prefix=None, n=None, devices=1,
             rebalance=1.25):
    """Headless sph_simulation::simulate(); returns dict(state, params, steps, frames[, recuts]).
    devices > 1: sph_simulation::devices — slabs over GPUs device .. device + devices - 1."""
    L = lib()
    if n is None:
        n = int(load_settings(fluid_json, sim_json)[0]["particles_count"][0])
    state = np.zeros(n, PARTICLE)
    params = np.zeros(1, PARAMS)
    steps = C.c_uint64()
    frames = C.c_uint32()
    err = C.create_string_buffer(512)
    if devices != 1:
        recuts = C.c_uint64()
        rc = L.clsph_simulate_devices(os.fsencode(fluid_json), os.fsencode(sim_json), os.fsencode(obj_path),
                                      max(device, 0), devices, float(rebalance), max_steps,
                                      os.fsencode(frames_prefix) if frames_prefix is not None else None,
                                      _ptr(state), _ptr(params), C.byref(steps), C.byref(frames),
                                      C.byref(recuts), err, len(err))
        if rc != 0:
            raise RuntimeError(err.value.decode())
        return dict(state=state, params=params, steps=steps.value, frames=frames.value, recuts=recuts.value)
    rc = L.clsph_simulate(os.fsencode(fluid_json), os.fsencode(sim_json), os.fsencode(obj_path), device,
                          max_steps, os.fsencode(frames_prefix) if frames_prefix is not None else None,
                          _ptr(state), _ptr(params), C.byref(steps), C.byref(frames), err, len(err))
    if rc != 0:
        raise RuntimeError(err.value.decode())
    return dict(state=state, params=params, steps=steps.value, frames=frames.value)
