"""Shared helpers of the test-suite (and of tests/golden/make_golden.py).

Nothing here reads /root/reference: property values are restated below, scenes are either
generated (axis-aligned open box) or come from the committed fixtures under tests/golden/.
"""
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
GOLDEN = os.path.join(ROOT, "tests", "golden")
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

# fluid_properties/water.json, fluid_properties/mucus.json, simulation_properties/default.json
WATER = dict(fluid_density=998.29, dynamic_viscosity=3.5, restitution=0, k=100,
             surface_tension_threshold=7.065, surface_tension=0.0728,
             particles_inside_influence_radius=20)
MUCUS = dict(fluid_density=1000, dynamic_viscosity=36, restitution=0.5, k=5,
             surface_tension_threshold=5, surface_tension=6,
             particles_inside_influence_radius=40)
DEFAULT_SIM = dict(particles_count=64000, particle_mass=0.05, simulation_time=3, target_fps=60,
                   simulation_scale=0.1, write_all_frames=False, serialize=False,
                   constant_acceleration=dict(x=0, y=-9.8, z=0))
FLUIDS = dict(water=WATER, mucus=MUCUS)

# the 10 triangles of an open-top box, in the vertex order / winding of scenes/box.obj
# (outward normals; the distance-field sign convention depends on them)
_BOX_FACES = [(1, 2, 3), (1, 5, 6), (2, 6, 3), (3, 7, 4), (5, 1, 4), (4, 1, 3), (2, 1, 6), (6, 7, 3),
              (7, 8, 4), (8, 5, 4)]


def box_mesh(xmin, xmax, ymin, ymax, zmin, zmax):
    """(vertices[8,3] f32, faces[10,3] int, 0-based) of an open-top axis-aligned box."""
    v = np.array([
        (xmax, ymin, zmin), (xmax, ymin, zmax), (xmin, ymin, zmax), (xmin, ymin, zmin),
        (xmax, ymax, zmin), (xmax, ymax, zmax), (xmin, ymax, zmax), (xmin, ymax, zmin)], np.float32)
    f = np.array(_BOX_FACES, np.int32) - 1
    return v, f


def write_obj(path, shapes):
    """shapes: list of (vertices[nv,3], faces[nf,3] 0-based within the shape)."""
    with open(path, "w") as fh:
        base = 0
        for k, (v, f) in enumerate(shapes):
            fh.write(f"o shape{k}\n")
            for p in np.asarray(v, np.float32):
                fh.write("v %.9g %.9g %.9g\n" % (p[0], p[1], p[2]))
            for t in np.asarray(f):
                fh.write("f %d %d %d\n" % (t[0] + 1 + base, t[1] + 1 + base, t[2] + 1 + base))
            base += len(v)


def dambreak_box(n):
    """The synthetic dam-break box of SURVEY.md 8(d) config 3, scaled to the lattice of n particles
    (water, m = 0.05): the block rests on the floor against the -x wall."""
    side = np.cbrt(n * 0.05 / 998.29)
    half = side / 2
    return box_mesh(-half - 0.06, half + 1.0 * side, -0.06, 2.2 * side, -half - 0.06, half + 0.06)


def rel_err(a, b, scale=None):
    """max |a-b| / scale, scale defaulting to max |b| (per array, not per component)."""
    a = np.asarray(a, np.float64)
    b = np.asarray(b, np.float64)
    if scale is None:
        scale = np.max(np.abs(b)) if b.size else 1.0
    if scale == 0:
        scale = 1.0
    return float(np.max(np.abs(a - b)) / scale) if a.size else 0.0


def xyz(arr):
    return np.ascontiguousarray(arr[:, :3])


def state_arrays(particles):
    """AoS structured array -> dict of plain arrays (xyz only for the float3 members)."""
    return dict(
        pos=xyz(particles["position"]), vel=xyz(particles["velocity"]),
        vhf=xyz(particles["intermediate_velocity"]), acc=xyz(particles["acceleration"]),
        rho=particles["density"].copy(), prs=particles["pressure"].copy(),
        key=particles["grid_index"].copy())


def write_json(path, d):
    import json
    with open(path, "w") as fh:
        fh.write(json.dumps(d, indent=4) + ";")  # the reference's property files end in '};'


def make_case(tmpdir, fluid="water", n=4096, box=None, particle_mass=None):
    """A synthetic case built with the PRODUCT's host code (libclsph_b200.host): settings from
    temporary JSON files, the reference lattice, and an open-top box scene (default: the
    dam-break box scaled to n).  Returns dict(params, terms, particles, scene)."""
    from libclsph_b200 import host
    tmpdir = str(tmpdir)
    sim = dict(DEFAULT_SIM, particles_count=int(n))
    if particle_mass is not None:
        sim["particle_mass"] = particle_mass
    write_json(os.path.join(tmpdir, "fluid.json"), FLUIDS[fluid])
    write_json(os.path.join(tmpdir, "sim.json"), sim)
    params, terms, vol, _, _ = host.load_settings(os.path.join(tmpdir, "fluid.json"), os.path.join(tmpdir, "sim.json"))
    if box is None:
        side = float(np.cbrt(vol))
        half = side / 2
        box = (-half - 0.06, half + 1.0 * side, -0.06, 2.2 * side, -half - 0.06, half + 0.06)
    v, f = box_mesh(*box)
    obj = os.path.join(tmpdir, "scene.obj")
    write_obj(obj, [(v, f)])
    sc = host.load_scene(obj, float(params["h"][0]) * 2)
    assert sc is not None
    return dict(params=params, terms=terms, particles=host.lattice(int(n), vol), scene=sc, volume=vol,
                fluid_json=os.path.join(tmpdir, "fluid.json"), sim_json=os.path.join(tmpdir, "sim.json"),
                obj=obj)


def new_engine(case, device=-1):
    from libclsph_b200 import Engine
    e = Engine(case["params"], case["terms"], device)
    sc = case["scene"]
    e.set_scene(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
    return e
