"""CPU: the host-side mirror of libclsph's API (libclsph_b200/host/, C++11) against fixtures
produced by the reference's own host code (tests/golden/make_golden.py): scene::load,
load_settings, the lattice of init_particles, and the Houdini .geo writer.  Also checks that
both product libraries load without a GPU and export every symbol their headers declare."""
import ctypes
import os
import re

import numpy as np
import pytest

from common import DEFAULT_SIM, FLUIDS, GOLDEN, ROOT, write_obj


@pytest.fixture(scope="session")
def built():
    from libclsph_b200 import build
    build.build_all()
    return True


def _declared(header):
    txt = open(os.path.join(ROOT, "include", header)).read()
    txt = re.sub(r"/\*.*?\*/", "", txt, flags=re.S)
    return sorted(set(re.findall(r"\b((?:sphb|clsph)_[a-z_0-9]+)\s*\(", txt)))


def test_sphb_exports_every_declared_symbol(built):
    from libclsph_b200 import engine
    L = engine.lib()
    names = _declared("sphb.h")
    assert len(names) >= 17
    for n in names:
        assert hasattr(L, n), n
    assert sorted(engine.SYMBOLS) == names


def test_host_exports_every_declared_symbol(built):
    from libclsph_b200 import host
    L = host.lib()
    names = [n for n in _declared("clsph_host.h") if n.startswith("clsph_")]
    for n in names:
        assert hasattr(L, n), n
    assert sorted(host.SYMBOLS) == names


def test_no_gpu_means_loud_failure(built):
    """There is no CPU path: without a device sphb_create fails with SPHB_ERR_NO_DEVICE."""
    import torch
    if torch.cuda.is_available():
        pytest.skip("a GPU is present")
    from libclsph_b200 import Engine, SphbError, PARAMS, TERMS
    p = np.zeros(1, PARAMS)
    p["particles_count"] = 64
    with pytest.raises(SphbError) as ei:
        Engine(p, np.zeros(1, TERMS))
    assert ei.value.code == -3


def test_struct_layouts(built):
    from libclsph_b200 import BB, PARAMS, PARTICLE, TERMS
    assert (PARTICLE.itemsize, PARAMS.itemsize, TERMS.itemsize, BB.itemsize) == (80, 128, 20, 56)
    assert PARTICLE.fields["density"][1] == 64 and PARTICLE.fields["grid_index"][1] == 72
    assert PARAMS.fields["constant_acceleration"][1] == 64 and PARAMS.fields["min_point"][1] == 96


@pytest.mark.parametrize("name", ["scene_box_water", "scene_cube_water", "scene_cone_water",
                                  "scene_dambreak_water", "scene_monkeybox_mucus", "scene_labyrinth_water",
                                  "scene_shower_water", "scene_monkey_water", "scene_plane_water",
                                  "scene_river_water"])
def test_scene_load_matches_reference(built, tmp_path, name):
    from libclsph_b200 import host
    z = np.load(os.path.join(GOLDEN, name + ".npz"))
    soup = z["soup"]
    v = soup.reshape(-1, 3)
    f = np.arange(len(v)).reshape(-1, 3)
    # two shapes, like monkeybox.obj: faces accumulate across shapes
    half = len(f) // 2
    shapes = [(v[: 3 * half], f[:half]), (v[3 * half:], f[half:] - 3 * half)] if half else [(v, f)]
    path = str(tmp_path / "scene.obj")
    write_obj(path, shapes)
    sc = host.load_scene(path, float(z["threshold"]))
    assert sc is not None
    assert sc["face_count"] == len(soup)
    assert sc["total_gridpoints"] == int(z["total_gridpoints"])
    for k in ("bbs", "transforms", "rvertices", "face_normals"):
        assert sc[k].tobytes() == z[k].tobytes(), k


def test_scene_load_errors(built, tmp_path):
    from libclsph_b200 import host
    assert host.load_scene(str(tmp_path / "missing.obj"), 0.1) is None
    # quads are fan-triangulated (tinyobj behaviour), negative indices are relative
    p = tmp_path / "quad.obj"
    p.write_text("v 0 0 0\nv 1 0 0\nv 1 0 1\nv 0 0 1\nf -4 -3 -2 -1\n")
    sc = host.load_scene(str(p), 0.1)
    assert sc["face_count"] == 2
    q = tmp_path / "tris.obj"
    q.write_text("v 0 0 0\nv 1 0 0\nv 1 0 1\nv 0 0 1\nf 1 2 3\nf 1 3 4\n")
    sc2 = host.load_scene(str(q), 0.1)
    for k in ("bbs", "transforms", "rvertices", "face_normals"):
        assert sc[k].tobytes() == sc2[k].tobytes()


def _write_json(path, d):
    import json
    path.write_text(json.dumps(d, indent=4) + ";")  # the reference's files end in '};'


@pytest.mark.parametrize("steps,fluid,n", [("steps_box_water", "water", 4096),
                                           ("steps_monkeybox_mucus", "mucus", 2744)])
def test_settings_and_lattice_match_reference(built, tmp_path, steps, fluid, n):
    from libclsph_b200 import host
    g = np.load(os.path.join(GOLDEN, steps + ".npz"))
    _write_json(tmp_path / "fluid.json", FLUIDS[fluid])
    _write_json(tmp_path / "sim.json", dict(DEFAULT_SIM, particles_count=n))
    p, t, vol, write_all, serialize = host.load_settings(str(tmp_path / "fluid.json"), str(tmp_path / "sim.json"))
    assert p.tobytes() == g["params0"].tobytes()
    assert t.tobytes() == g["terms"].tobytes()
    assert (write_all, serialize) == (False, False)
    lat = host.lattice(n, vol)
    assert lat.tobytes() == g["initial"].tobytes()


def test_settings_reject_bad_restitution(built, tmp_path):
    from libclsph_b200 import host
    _write_json(tmp_path / "fluid.json", dict(FLUIDS["water"], restitution=1.5))
    _write_json(tmp_path / "sim.json", DEFAULT_SIM)
    with pytest.raises(RuntimeError, match="Restitution has an invalid value!"):
        host.load_settings(str(tmp_path / "fluid.json"), str(tmp_path / "sim.json"))


def test_geo_writer_matches_reference(built):
    from libclsph_b200 import host
    z = np.load(os.path.join(GOLDEN, "geo_frame.npz"))
    ref = z["text"].tobytes()
    for threads in (1, 3):
        assert host.geo_text(z["particles"], float(z["particle_mass"]), threads) == ref
    # the parallel path (>= 4096 particles) must give the same bytes as the serial one
    big = np.tile(z["particles"], 20)
    assert host.geo_text(big, 0.05, 1) == host.geo_text(big, 0.05, 7)
