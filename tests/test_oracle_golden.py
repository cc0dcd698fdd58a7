"""CPU: pins the plain-C oracle (oracle/sph_oracle.c) against fixtures produced by the
reference's own kernel source (tests/golden/make_golden.py).  Bit-exact everywhere: both are
compiled with -ffp-contract=off and evaluate the same expression trees."""
import hashlib
import os

import numpy as np
import pytest

from common import GOLDEN, state_arrays

STEP_FIXTURES = [("steps_box_water", "scene_box_water"), ("steps_dambreak_water", "scene_dambreak_water"),
                 ("steps_monkeybox_mucus", "scene_monkeybox_mucus"),
                 # a block thrown INTO the collision-heavy meshes (make_golden.py: collision_steps)
                 ("steps_labyrinth_water", "scene_labyrinth_water"), ("steps_river_water", "scene_river_water")]


def load_scene(name):
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    return z


def bits(a):
    a = np.ascontiguousarray(a)
    return a.view(np.uint32) if a.dtype == np.float32 else a


def test_struct_sizes(port_oracle):
    assert port_oracle.kind == "port"


ALL_SCENES = ["scene_box_water", "scene_cube_water", "scene_cone_water",
                                   "scene_dambreak_water", "scene_monkeybox_mucus", "scene_labyrinth_water",
                                   "scene_shower_water", "scene_monkey_water", "scene_plane_water",
                                   "scene_river_water"]


def df_sample(df):
    """the fixture's sampling of a distance field (tests/golden/make_golden.py)"""
    return df[::97] if len(df) < (1 << 21) else df[::997]


@pytest.mark.parametrize("scene", ALL_SCENES)
def test_distance_field_matches_reference(port_oracle, scene):
    z = load_scene(scene)
    df = port_oracle.compute_df(z["bbs"], z["transforms"], z["rvertices"], int(z["total_gridpoints"]))
    assert np.array_equal(bits(df_sample(df)), bits(z["df_sample"]))
    assert hashlib.sha256(df.tobytes()).hexdigest() == str(z["df_sha256"])


@pytest.mark.parametrize("steps,scene", STEP_FIXTURES)
def test_steps_match_reference(port_oracle, steps, scene):
    g = np.load(os.path.join(GOLDEN, steps + ".npz"))
    sc = load_scene(scene)
    orc = port_oracle
    df = orc.compute_df(sc["bbs"], sc["transforms"], sc["rvertices"], int(sc["total_gridpoints"]))
    params = g["params0"].copy()
    terms = g["terms"].copy()
    front = g["initial"].copy()
    dt = float(g["dt0"])
    for s in range(int(g["nsteps"])):
        r = orc.step(front, params, terms, df, sc["bbs"], dt)
        dt = r["dt"]
        assert np.float32(dt).view(np.uint32) == g[f"s{s}_dt"].view(np.uint32)
        assert r["attempts"] == int(g[f"s{s}_attempts"])
        assert np.array_equal(r["perm"], g[f"s{s}_perm"])
        assert np.array_equal(r["cell_table"], g[f"s{s}_cell_table"])
        assert params.tobytes() == g[f"s{s}_params"].tobytes()
        st = state_arrays(front)
        for k, v in st.items():
            assert np.array_equal(bits(v), bits(g[f"s{s}_{k}"])), (s, k)


@pytest.mark.parametrize("steps", ["steps_labyrinth_water", "steps_river_water"])
def test_collision_fixtures_do_collide_and_retry(steps):
    """The in-mesh fixtures are worth having only if handle_collisions fires DURING the run
    (kernels/collisionsv2.cl:249-336) and the advection is retried (sph_simulation.cpp:246-260)."""
    g = np.load(os.path.join(GOLDEN, steps + ".npz"))
    attempts = [int(g[f"s{s}_attempts"]) for s in range(int(g["nsteps"]))]
    assert int(g["nsteps"]) >= 10 and max(attempts) > 1
    collided = []
    vh = g["initial"]["intermediate_velocity"][:, :3]
    for s in range(int(g["nsteps"])):
        free = vh[g[f"s{s}_perm"]] + g[f"s{s}_acc"] * g[f"s{s}_dt"]  # v' = v_half + a dt without a response
        if attempts[s] == 1:  # (after a retry the dt of the last attempt is not the step's output dt)
            dt_used = np.float32(g["dt0"]) if s == 0 else g[f"s{s-1}_dt"]
            free = vh[g[f"s{s}_perm"]] + g[f"s{s}_acc"] * np.float32(dt_used)
            collided.append(int(np.any(free != g[f"s{s}_vhf"], axis=1).sum()))
        vh = g[f"s{s}_vhf"]
    assert len(collided) >= 8 and min(collided) > 200, collided


def test_settings_match_reference(port_oracle):
    from common import DEFAULT_SIM, FLUIDS
    for steps, fluid, n in [("steps_box_water", "water", 4096), ("steps_monkeybox_mucus", "mucus", 2744)]:
        g = np.load(os.path.join(GOLDEN, steps + ".npz"))
        p, t, vol = port_oracle.settings(FLUIDS[fluid], DEFAULT_SIM, particles_count=n)
        assert p.tobytes() == g["params0"].tobytes()
        assert t.tobytes() == g["terms"].tobytes()
        init = port_oracle.init_particles(p, vol)
        assert init.tobytes() == g["initial"].tobytes()


def test_advect_random_matches_reference(port_oracle):
    g = np.load(os.path.join(GOLDEN, "advect_random.npz"))
    sc = load_scene("scene_monkeybox_mucus")
    df = port_oracle.compute_df(sc["bbs"], sc["transforms"], sc["rvertices"], int(sc["total_gridpoints"]))
    out = port_oracle.advect(g["input"], float(g["restitution"]), float(g["dt"]), df, sc["bbs"])
    for f in ("position", "velocity", "intermediate_velocity"):
        a, b = out[f][:, :3], g["output"][f][:, :3]
        same = (bits(a) == bits(b)) | (np.isnan(a) & np.isnan(b))
        assert same.all(), f


def test_port_equals_reference_live(port_oracle, ref_oracle):
    """When oracle/_ref is present (it travels to the GPU box as a built .so): one more config,
    not in the fixtures, run through both checkers."""
    from common import DEFAULT_SIM, FLUIDS
    sc = load_scene("scene_dambreak_water")
    outs = []
    for orc in (port_oracle, ref_oracle):
        p, t, vol = orc.settings(FLUIDS["water"], DEFAULT_SIM, particles_count=3375)
        df = orc.compute_df(sc["bbs"], sc["transforms"], sc["rvertices"], int(sc["total_gridpoints"]))
        front = orc.init_particles(p, vol)
        dt = 1.0 / 600.0
        for _ in range(4):
            dt = orc.step(front, p, t, df, sc["bbs"], dt)["dt"]
        outs.append((front.copy(), dt))
    assert outs[0][1] == outs[1][1]
    for f in ("position", "velocity", "intermediate_velocity", "acceleration", "density", "pressure", "grid_index"):
        assert np.array_equal(bits(outs[0][0][f]), bits(outs[1][0][f])), f
