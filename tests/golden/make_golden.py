"""Generates the committed fixtures under tests/golden/ from the REFERENCE ITSELF.

Run in the build container only (needs /root/reference and oracle/_ref/libsphref.so, the
reference's own kernels/sph.cl + scene.cpp compiled as C++, see oracle/Makefile):

    python tests/golden/make_golden.py

Fixtures (all produced by the "reference" checker, never by the port or the CUDA engine):
  scene_<name>.npz   scene::load products (bbs, transforms, rvertices, normals, total grid
                     points) of a reference scene at a given distFieldThreshold, the triangle
                     soup they came from, and a strided sample + checksum of the distance field
  steps_<name>.npz   k successive simulate_single_frame results from the reference lattice:
                     per step keys, permutation, cell table, rho, p, a, x, v, v_half, dt, attempts
                     (steps_labyrinth_water / steps_river_water: a block thrown at the mesh —
                     collisions on every step, advection retries on the first ones)
  advect_random.npz  one advection_collision over a seeded random state inside monkeybox
"""
import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, os.path.dirname(HERE))
from common import DEFAULT_SIM, FLUIDS, ROOT, box_mesh, dambreak_box, state_arrays, write_obj  # noqa: E402

from oracle import pyorc  # noqa: E402

REF = "/root/reference"


def parse_obj_triangles(path):
    """Triangle soup [nf,3,3] f32 in file order (fan triangulation, like tinyobj)."""
    verts, tris = [], []
    for line in open(path):
        t = line.split()
        if not t:
            continue
        if t[0] == "v":
            verts.append([np.float32(x) for x in t[1:4]])
        elif t[0] == "f":
            idx = [int(tok.split("/")[0]) for tok in t[1:]]
            idx = [i - 1 if i > 0 else len(verts) + i for i in idx]
            for k in range(1, len(idx) - 1):
                tris.append([verts[idx[0]], verts[idx[k]], verts[idx[k + 1]]])
    return np.array(tris, np.float32)


def scene_fixture(orc, name, root_dir, filename, threshold, soup, out):
    sc = orc.ref_scene(root_dir, filename, threshold)
    assert sc is not None, filename
    df = orc.compute_df(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
    np.savez_compressed(
        out, threshold=np.float32(threshold), soup=soup, bbs=sc["bbs"], transforms=sc["transforms"],
        rvertices=sc["rvertices"], face_normals=sc["face_normals"],
        total_gridpoints=np.uint32(sc["total_gridpoints"]),
        df_sample=df[::97].copy() if len(df) < (1 << 21) else df[::997].copy(),
        df_sha256=np.array(hashlib.sha256(df.tobytes()).hexdigest()))
    print(name, "faces", sc["face_count"], "grid points", sc["total_gridpoints"])
    return sc, df


def steps_fixture(orc, fluid, n, sc, df, nsteps, out, prepare=None):
    params, terms, vol = orc.settings(FLUIDS[fluid], DEFAULT_SIM, particles_count=n)
    front = orc.init_particles(params, vol)
    if prepare:
        prepare(front)
    dt = np.float32(1.0) / params["target_fps"][0] * params["simulation_scale"][0]
    rec = dict(n=np.uint32(n), fluid=np.array(fluid), initial=front.copy(), params0=params.copy(),
               terms=terms.copy(), dt0=np.float32(dt))
    for s in range(nsteps):
        r = orc.step(front, params, terms, df, sc["bbs"], float(dt))
        dt = r["dt"]
        st = state_arrays(front)
        for k, v in st.items():
            rec[f"s{s}_{k}"] = v
        rec[f"s{s}_perm"] = r["perm"]
        rec[f"s{s}_cell_table"] = r["cell_table"]
        rec[f"s{s}_dt"] = np.float32(r["dt"])
        rec[f"s{s}_attempts"] = np.int32(r["attempts"])
        rec[f"s{s}_params"] = params.copy()
        print("  step", s, "dt", r["dt"], "attempts", r["attempts"], "grid",
              params["grid_size_x"][0], params["grid_size_y"][0], params["grid_size_z"][0])
    rec["nsteps"] = np.int32(nsteps)
    np.savez_compressed(out, **rec)


def main():
    pyorc.build("reference")
    orc = pyorc.load("reference")
    hw = float(orc.settings(FLUIDS["water"], DEFAULT_SIM)[0]["h"][0])
    hm = float(orc.settings(FLUIDS["mucus"], DEFAULT_SIM)[0]["h"][0])

    # reference scenes
    scenes = {}
    for name, fn, h in [("box_water", "box.obj", hw), ("monkeybox_mucus", "monkeybox.obj", hm),
                        ("cube_water", "cube.obj", hw), ("cone_water", "cone.obj", hw)]:
        soup = parse_obj_triangles(os.path.join(REF, "scenes", fn))
        scenes[name] = scene_fixture(orc, name, REF, fn, np.float32(h) * 2, soup,
                                     os.path.join(HERE, f"scene_{name}.npz"))

    # synthetic dam-break box (generated OBJ, loaded by the reference's scene.cpp)
    n_dam = 4096
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "scenes"))
        v, f = dambreak_box(n_dam)
        write_obj(os.path.join(tmp, "scenes", "dam.obj"), [(v, f)])
        soup = v[f]
        scenes["dambreak_water"] = scene_fixture(orc, "dambreak_water", tmp, "dam.obj", np.float32(hw) * 2,
                                                 soup, os.path.join(HERE, "scene_dambreak_water.npz"))

    print("steps box_water")
    steps_fixture(orc, "water", 4096, *scenes["box_water"], 3, os.path.join(HERE, "steps_box_water.npz"))
    print("steps dambreak_water")
    steps_fixture(orc, "water", n_dam, *scenes["dambreak_water"], 6, os.path.join(HERE, "steps_dambreak_water.npz"))
    print("steps monkeybox_mucus")
    steps_fixture(orc, "mucus", 2744, *scenes["monkeybox_mucus"], 3,
                  os.path.join(HERE, "steps_monkeybox_mucus.npz"))

    # advection + collision over a random state inside monkeybox (mucus: restitution 0.5)
    sc, df = scenes["monkeybox_mucus"]
    rng = np.random.default_rng(20261104)
    n = 8192
    p = np.zeros(n, pyorc.PARTICLE)
    lo = np.array([-2.1, -2.1, -2.1], np.float32)
    hi = np.array([2.1, 3.0, 2.1], np.float32)
    p["position"][:, :3] = rng.uniform(lo, hi, (n, 3)).astype(np.float32)
    p["intermediate_velocity"][:, :3] = rng.normal(0, 2.0, (n, 3)).astype(np.float32)
    p["velocity"][:, :3] = p["intermediate_velocity"][:, :3]
    p["acceleration"][:, :3] = rng.normal(0, 30.0, (n, 3)).astype(np.float32)
    p["intermediate_velocity"][:16, :3] = 0  # |v| = 0 -> the reference's 0/0 in respond()
    p["acceleration"][:16, :3] = 0
    dt = np.float32(1.0 / 600.0)
    out = orc.advect(p, 0.5, float(dt), df, sc["bbs"])
    np.savez_compressed(os.path.join(HERE, "advect_random.npz"), input=p, output=out, dt=dt,
                        restitution=np.float32(0.5))
    moved = np.any(out["position"][:, :3] != (p["position"][:, :3] + (p["intermediate_velocity"][:, :3] + p["acceleration"][:, :3] * dt) * dt), axis=1)
    print("advect_random: collided", int(moved.sum()), "of", n)


def extra_scenes(orc):
    """The reference's remaining scenes at water's h (scene::load products + distance-field
    checksum): the collision-heavy meshes of SURVEY.md 8(d) config 5 and the small ones."""
    hw = float(orc.settings(FLUIDS["water"], DEFAULT_SIM)[0]["h"][0])
    for name, fn in [("labyrinth_water", "labyrinth.obj"), ("river_water", "river.obj"),
                     ("shower_water", "shower.obj"), ("monkey_water", "monkey.obj"), ("plane_water", "plane.obj")]:
        soup = parse_obj_triangles(os.path.join(REF, "scenes", fn))
        scene_fixture(orc, name, REF, fn, np.float32(hw) * 2, soup, os.path.join(HERE, f"scene_{name}.npz"))


def collision_steps(orc):
    """Step sequences INSIDE the collision-heavy meshes of SURVEY.md 8(d) config 5 (labyrinth.obj,
    river.obj): the reference lattice of 2 744 water particles is moved to just above a face near
    the middle of the mesh and thrown at it (seeded velocities, ~60 m/s downwards), so that
    handle_collisions (kernels/collisionsv2.cl:249-336) fires for ~1/3 of the particles on every
    step and the advection retry of sph_simulation.cpp:246-260 (attempts > 1) fires on the first
    steps.  12 steps each."""
    hw = float(orc.settings(FLUIDS["water"], DEFAULT_SIM)[0]["h"][0])
    for name, fn, face in [("labyrinth_water", "labyrinth.obj", 81), ("river_water", "river.obj", 85)]:
        soup = parse_obj_triangles(os.path.join(REF, "scenes", fn))
        sc = orc.ref_scene(REF, fn, np.float32(hw) * 2)
        df = orc.compute_df(sc["bbs"], sc["transforms"], sc["rvertices"], sc["total_gridpoints"])
        c = soup.mean(axis=1)[face]

        def prepare(front, c=c):
            n = len(front)
            rng = np.random.default_rng(5)
            front["position"][:, 0] += np.float32(c[0])
            front["position"][:, 1] += np.float32(c[1] + 0.03)
            front["position"][:, 2] += np.float32(c[2])
            v = np.zeros((n, 3), np.float32)
            v[:, 1] = (-60 + rng.normal(0, 5, n)).astype(np.float32)
            v[:, 0] = rng.normal(0, 3, n).astype(np.float32)
            v[:, 2] = rng.normal(0, 3, n).astype(np.float32)
            front["velocity"][:, :3] = v
            front["intermediate_velocity"][:, :3] = v

        print("steps", name)
        steps_fixture(orc, "water", 2744, sc, df, 12, os.path.join(HERE, f"steps_{name}.npz"), prepare)


def geo_fixture(orc):
    """One frame written by the reference's own houdini_file_saver (+ HoudiniFileDumpHelper)."""
    import ctypes as C
    rng = np.random.default_rng(7)
    n = 257
    p = np.zeros(n, pyorc.PARTICLE)
    p["position"][:, :3] = rng.normal(0, 3.0, (n, 3)).astype(np.float32)
    p["velocity"][:, :3] = (rng.normal(0, 1.0, (n, 3)) * 10.0 ** rng.integers(-8, 6, (n, 1))).astype(np.float32)
    p["density"] = rng.uniform(-50, 2200, n).astype(np.float32)
    p["density"][:6] = [0, 500, 1000, 1500, 2000, 999.99994]
    p["position"][0, :3] = [0, -0.0, 1e-7]
    p["position"][1, :3] = [123456.7, 1234567.0, 1e10]
    p["velocity"][2, :3] = [np.nan, np.inf, -np.inf]
    params, _, _ = orc.settings(FLUIDS["water"], DEFAULT_SIM, particles_count=n)
    orc.lib.orcgeo_write.argtypes = [C.c_char_p, C.c_void_p, C.c_void_p, C.c_int]
    with tempfile.TemporaryDirectory() as tmp:
        os.makedirs(os.path.join(tmp, "frames"))
        orc.lib.orcgeo_write((tmp + "/").encode(), p.ctypes.data, params.ctypes.data, 2)
        names = sorted(os.listdir(os.path.join(tmp, "frames")))
        text = open(os.path.join(tmp, "frames", names[0]), "rb").read()
    np.savez_compressed(os.path.join(HERE, "geo_frame.npz"), particles=p, particle_mass=params["particle_mass"][0],
                        names=np.array(names), text=np.frombuffer(text, np.uint8))
    print("geo_frame:", names, len(text), "bytes")


if __name__ == "__main__":
    if "--extra-scenes" in sys.argv:
        pyorc.build("reference")
        extra_scenes(pyorc.load("reference"))
    elif "--collision-steps" in sys.argv:
        pyorc.build("reference")
        collision_steps(pyorc.load("reference"))
    elif "--geo-only" in sys.argv:
        pyorc.build("reference")
        geo_fixture(pyorc.load("reference"))
    else:
        main()
        extra_scenes(pyorc.load("reference"))
        collision_steps(pyorc.load("reference"))
        geo_fixture(pyorc.load("reference"))
