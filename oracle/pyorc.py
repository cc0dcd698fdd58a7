"""ctypes/numpy front-end of the two CPU checkers (oracle/oracle_api.h).

TEST INFRASTRUCTURE ONLY: imported by tests/, by __graft_entry__.smoke() and by
bench.py's CPU-baseline legs.  The product package (libclsph_b200/) never imports
this module.

    orc = pyorc.load("port")        # oracle/liboracle.so   (oracle/sph_oracle.c)
    orc = pyorc.load("reference")   # oracle/_ref/libsphref.so (the reference's sph.cl as C++)
"""
import ctypes as C
import json
import os
import re
import subprocess

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))

f3 = np.dtype((np.float32, 4))
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
    ],
    align=False,
)
assert PARTICLE.itemsize == 80

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
    ],
    align=False,
)
assert PARAMS.itemsize == 128

TERMS = np.dtype(
    [
        ("poly_6", np.float32),
        ("poly_6_gradient", np.float32),
        ("poly_6_laplacian", np.float32),
        ("spiky", np.float32),
        ("viscosity", np.float32),
    ]
)
assert TERMS.itemsize == 20

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
    ],
    align=False,
)
assert BB.itemsize == 56


def _ptr(a):
    return a.ctypes.data_as(C.c_void_p) if a is not None else None


def lib_path(kind):
    if kind == "reference":
        return os.path.join(HERE, "_ref", "libsphref.so")
    if kind == "port":  # SPHB_ORACLE_LIB: another build of it (tools/sanitize_cpu.sh)
        return os.environ.get("SPHB_ORACLE_LIB") or os.path.join(HERE, "liboracle.so")
    raise ValueError(kind)


def build(kind):
    """(Re)build a checker with oracle/Makefile.  'reference' needs /root/reference."""
    subprocess.check_call(["make", "-s", "-C", HERE, kind])


def available(kind):
    return os.path.exists(lib_path(kind))


def read_json(path):
    """The reference's property files end in '};' — picojson stops after the value."""
    txt = open(path).read()
    txt = txt[: txt.rindex("}") + 1]
    return json.loads(txt)


class Oracle:
    def __init__(self, kind):
        self.kind = kind
        self.lib = C.CDLL(lib_path(kind))
        L = self.lib
        L.orc_kind.restype = C.c_char_p
        L.orc_get_threads.restype = C.c_int
        L.orc_find_min_max.restype = C.c_int
        L.orc_compute_dt.restype = C.c_float
        L.orc_step.restype = C.c_float
        L.orc_step.argtypes = [C.c_void_p] * 6 + [C.c_uint32, C.c_float, C.c_void_p, C.c_void_p,
                                                  C.c_uint32, C.c_void_p]
        L.orc_advect.argtypes = [C.c_void_p, C.c_void_p, C.c_uint32, C.c_float, C.c_float,
                                 C.c_void_p, C.c_void_p, C.c_uint32]
        L.orc_sort.argtypes = [C.c_void_p] * 5 + [C.c_int]
        L.orc_density.argtypes = [C.c_void_p] * 5
        L.orc_forces.argtypes = [C.c_void_p] * 5
        L.orc_locate.argtypes = [C.c_void_p] * 3
        L.orc_find_min_max.argtypes = [C.c_void_p] * 2
        L.orc_compute_dt.argtypes = [C.c_void_p] * 2
        L.orc_compute_df.argtypes = [C.c_void_p] * 4 + [C.c_uint32, C.c_uint32]
        L.orc_derive.argtypes = [C.c_void_p, C.c_void_p, C.c_int, C.c_void_p]
        L.orc_init_particles.argtypes = [C.c_void_p, C.c_void_p, C.c_float]
        L.orc_set_threads.argtypes = [C.c_int]
        assert L.orc_kind().decode() == kind
        sizes = (C.c_uint32 * 4)()
        L.orc_sizes(sizes)
        assert list(sizes) == [80, 128, 20, 56], list(sizes)
        # every host core this process may run on — NOT omp_get_max_threads(), which follows
        # OMP_NUM_THREADS, and torch.distributed.run sets that to 1 for every rank
        try:
            self.cores = len(os.sched_getaffinity(0))
        except AttributeError:
            self.cores = os.cpu_count() or 1
        L.orc_set_threads(self.cores)

    # ---- settings -------------------------------------------------------------------
    def settings(self, fluid, sim, particles_count=None, particle_mass=None):
        """Restates sph_simulation::load_settings (sph_simulation.cpp:434-532).
        fluid/sim: dicts or paths of the two JSON files.  Returns (params, terms, initial_volume)."""
        if isinstance(fluid, str):
            fluid = read_json(fluid)
        if isinstance(sim, str):
            sim = read_json(sim)
        p = np.zeros(1, PARAMS)
        t = np.zeros(1, TERMS)
        p["fluid_density"] = fluid["fluid_density"]
        p["dynamic_viscosity"] = fluid["dynamic_viscosity"]
        p["restitution"] = fluid["restitution"]
        if not (0 <= p["restitution"][0] <= 1):
            raise RuntimeError("Restitution has an invalid value!")
        p["K"] = fluid["k"]
        p["surface_tension_threshold"] = fluid["surface_tension_threshold"]
        p["surface_tension"] = fluid["surface_tension"]
        n_infl = int(fluid["particles_inside_influence_radius"])
        p["particles_count"] = int(sim["particles_count"] if particles_count is None else particles_count)
        p["particle_mass"] = sim["particle_mass"] if particle_mass is None else particle_mass
        p["simulation_time"] = sim["simulation_time"]
        p["target_fps"] = sim["target_fps"]
        p["simulation_scale"] = sim["simulation_scale"]
        ca = sim["constant_acceleration"]
        p["constant_acceleration"][0, :3] = [ca["x"], ca["y"], ca["z"]]
        vol = C.c_float()
        self.lib.orc_derive(_ptr(p), _ptr(t), n_infl, C.byref(vol))
        return p, t, vol.value

    def init_particles(self, params, initial_volume):
        n = int(params["particles_count"][0])
        buf = np.zeros(n, PARTICLE)
        self.lib.orc_init_particles(_ptr(buf), _ptr(params), initial_volume)
        return buf

    # ---- passes ---------------------------------------------------------------------
    def find_min_max(self, particles, params):
        return self.lib.orc_find_min_max(_ptr(particles), _ptr(params))

    def locate(self, particles, params):
        out = np.zeros_like(particles)
        self.lib.orc_locate(_ptr(particles), _ptr(out), _ptr(params))
        return out

    def sort(self, particles, params, thread_count=0):
        """returns (sorted, perm, cell_table)"""
        first = particles.copy()
        second = np.zeros_like(first)
        perm = np.zeros(len(first), np.uint32)
        table = np.zeros(int(params["grid_cell_count"][0]), np.uint32)
        self.lib.orc_sort(_ptr(first), _ptr(second), _ptr(table), _ptr(params), _ptr(perm), thread_count)
        return first, perm, table

    def density(self, particles, params, terms, cell_table):
        out = np.zeros_like(particles)
        self.lib.orc_density(_ptr(particles), _ptr(out), _ptr(params), _ptr(terms), _ptr(cell_table))
        return out

    def forces(self, particles, params, terms, cell_table):
        out = np.zeros_like(particles)
        self.lib.orc_forces(_ptr(particles), _ptr(out), _ptr(params), _ptr(terms), _ptr(cell_table))
        return out

    def advect(self, particles, restitution, dt, df, bbs):
        out = np.zeros_like(particles)
        self.lib.orc_advect(_ptr(particles), _ptr(out), len(particles), restitution, dt,
                            _ptr(df), _ptr(bbs), len(bbs))
        return out

    def compute_dt(self, particles, params):
        return self.lib.orc_compute_dt(_ptr(particles), _ptr(params))

    def compute_df(self, bbs, transforms, rvertices, gridcount):
        df = np.zeros(gridcount, np.float32)
        self.lib.orc_compute_df(_ptr(df), _ptr(bbs), _ptr(transforms), _ptr(rvertices), len(bbs), gridcount)
        return df

    def step(self, front, params, terms, df, bbs, dt, cell_table_cap=1 << 22):
        """One simulate_single_frame in place on `front` (and params).
        returns dict(dt, perm, cell_table, attempts)"""
        back = np.zeros_like(front)
        perm = np.zeros(len(front), np.uint32)
        table = np.zeros(cell_table_cap, np.uint32)
        attempts = C.c_int()
        out = self.lib.orc_step(_ptr(front), _ptr(back), _ptr(params), _ptr(terms), _ptr(df),
                                _ptr(bbs), len(bbs), dt, _ptr(perm), _ptr(table), cell_table_cap,
                                C.byref(attempts))
        ncell = int(params["grid_cell_count"][0])
        return dict(dt=out, perm=perm, cell_table=table[:ncell].copy(), attempts=attempts.value)

    def last_step_times(self):
        t = (C.c_double * 6)()
        self.lib.orc_last_step_times(t)
        return dict(zip(["minmax", "locate", "sort", "density", "forces", "advect"], list(t)))

    # ---- reference-only: the reference's scene.cpp ------------------------------------
    def ref_scene(self, root_dir, filename, threshold):
        """Runs the reference's scene::load on <root_dir>/scenes/<filename>.  'reference' only."""
        assert self.kind == "reference"
        L = self.lib
        L.orcscene_load.restype = C.c_void_p
        L.orcscene_load.argtypes = [C.c_char_p, C.c_char_p, C.c_float]
        L.orcscene_face_count.argtypes = [C.c_void_p]
        L.orcscene_total_gridpoints.argtypes = [C.c_void_p]
        L.orcscene_copy.argtypes = [C.c_void_p] * 5
        L.orcscene_free.argtypes = [C.c_void_p]
        h = L.orcscene_load(root_dir.encode(), filename.encode(), threshold)
        if not h:
            return None
        nf = L.orcscene_face_count(h)
        sc = dict(
            face_count=nf,
            total_gridpoints=L.orcscene_total_gridpoints(h),
            bbs=np.zeros(nf, BB),
            transforms=np.zeros(nf * 12, np.float32),
            rvertices=np.zeros(nf * 4, np.float32),
            face_normals=np.zeros(nf * 3, np.float32),
        )
        L.orcscene_copy(h, _ptr(sc["bbs"]), _ptr(sc["transforms"]), _ptr(sc["rvertices"]),
                        _ptr(sc["face_normals"]))
        L.orcscene_free(h)
        return sc


_cache = {}


def load(kind="port"):
    if kind not in _cache:
        _cache[kind] = Oracle(kind)
    return _cache[kind]
