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


def _bgeo_expected(particles, mass, h):
    """Independent restatement (numpy + struct) of what partio's Bgeo-V5 writer emits for the
    attributes makeData adds (util/partio/PartioFunctions.h:5-65): header, five point
    attributes, points, the one particle-system primitive, trailer.  partio is not vendored in
    the reference and not installed here: parity of the .bgeo branch is UNPINNED (this pins the
    C++ writer to the format restated here, not to bytes a partio build produced)."""
    import struct
    n = len(particles)
    hs = lambda s: struct.pack(">H", len(s)) + s
    out = struct.pack(">IcIIIIIIIII", 0x4267656F, b"V", 5, n, 1, 0, 0, 5, 0, 1, 0)
    for name, size, typ in ((b"velocity", 3, 5), (b"color", 3, 5), (b"id", 1, 1), (b"mass", 1, 0),
                            (b"pscale", 1, 0)):
        out += hs(name) + struct.pack(">Hi", size, typ) + b"\0" * (4 * size)
    d = particles["density"].astype(np.float32)
    one, k, half = np.float32(1), np.float32(1000), np.float32(500)
    with np.errstate(invalid="ignore"):
        r = np.where((d > k) & (d <= 2 * k), (d - k) / k, np.float32(0))
        g = np.where((d >= 0) & (d < k), one - d / k, np.float32(0))
        b = np.where((d >= half) & (d <= k), (d - half) / half,
                     np.where((d >= k) & (d <= np.float32(1500)), one - (d - k) / half, np.float32(0)))
    rec = np.zeros(n, dtype=[("p", ">f4", 4), ("v", ">f4", 3), ("c", ">f4", 3), ("id", ">i4"), ("m", ">f4"),
                             ("s", ">f4")])
    rec["p"][:, :3] = particles["position"][:, :3]
    rec["p"][:, 3] = 1.0
    rec["v"][:, 0] = particles["velocity"][:, 0]  # y, z never written by makeData (:36-38)
    rec["c"] = np.stack([r, g, b], axis=1).astype(np.float32)
    rec["id"] = np.arange(n)
    rec["m"] = mass
    rec["s"] = h
    out += rec.tobytes()
    out += hs(b"generator") + struct.pack(">Hii", 1, 4, 1) + hs(b"papi")
    out += struct.pack(">ii", 0x8000, n)
    out += np.arange(n).astype(">u4" if n > 65536 else ">u2").tobytes()
    out += struct.pack(">i", 0) + b"\x00\xff"
    return out


@pytest.mark.parametrize("count", [0, 1, 100, 5000, 70000])
def test_bgeo_writer_matches_format(built, count):
    from libclsph_b200 import host
    from libclsph_b200.engine import PARAMS
    z = np.load(os.path.join(GOLDEN, "geo_frame.npz"))
    rng = np.random.default_rng(7)
    ps = np.tile(z["particles"], count // len(z["particles"]) + 1)[:count].copy()
    if count:
        ps["density"] = rng.uniform(-100, 2300, count).astype(np.float32)
        ps["density"][::17] = np.float32(1000)
        ps["density"][::19] = np.nan
    p = np.zeros(1, PARAMS)
    p["particles_count"] = count
    p["particle_mass"] = 0.05
    p["h"] = 0.06207
    want = _bgeo_expected(ps, np.float32(0.05), np.float32(0.06207))
    for threads in (1, 5):
        assert host.bgeo_bytes(ps, p, threads) == want


def test_saver_writes_geo_or_bgeo(built, tmp_path):
    from libclsph_b200 import host
    from libclsph_b200.engine import PARAMS
    z = np.load(os.path.join(GOLDEN, "geo_frame.npz"))
    ps = z["particles"]
    p = np.zeros(1, PARAMS)
    p["particles_count"] = len(ps)
    p["particle_mass"] = float(z["particle_mass"])
    p["h"] = 0.1
    (tmp_path / "frames").mkdir()
    prefix = str(tmp_path) + "/"
    host.save_frame(prefix, ps, p, use_partio=False)
    host.save_frame(prefix, ps, p, use_partio=True)
    assert (tmp_path / "frames" / "frame0000001.geo").read_bytes() == z["text"].tobytes()
    assert (tmp_path / "frames" / "frame0000001.bgeo").read_bytes() == host.bgeo_bytes(ps, p)


@pytest.mark.parametrize("world", [1, 2, 3, 8])
def test_slab_faces_equal_counts(built, world):
    """The faces sph_simulation::devices deals the particles out by (host/sph_simulation_slabs.cpp)
    are those of the Python driver (libclsph_b200/slab.py): equal counts, a lattice plane never
    split, -inf / +inf outside; also on a cloud of distinct random x values."""
    from libclsph_b200 import PARTICLE, host
    from libclsph_b200.slab import equal_count_cuts
    rng = np.random.default_rng(7)
    lattice = host.lattice(4096, 0.2)
    cloud = np.zeros(5000, PARTICLE)
    cloud["position"][:, 0] = rng.normal(size=5000).astype(np.float32)
    few = cloud[:3].copy()  # fewer particles than slabs: faces may coincide, nothing may be lost
    got = host.equal_count_cuts(few, world)
    want = np.array(equal_count_cuts(few["position"][:, 0], world), np.float32)
    assert np.array_equal(got, want)
    for p in (lattice, cloud):
        got = host.equal_count_cuts(p, world)
        want = np.array(equal_count_cuts(p["position"][:, 0], world), np.float32)
        assert np.array_equal(got, want)
        assert got[0] == -np.inf and got[-1] == np.inf and np.all(np.diff(got) > 0)
        x = p["position"][:, 0]
        counts = [int(((x >= got[r]) & (x < got[r + 1])).sum()) for r in range(world)]
        assert sum(counts) == len(p)
        if p is cloud:
            assert max(counts) - min(counts) <= 1
        else:  # whole lattice planes of 256 particles each
            assert all(c % 256 == 0 for c in counts) and max(counts) - min(counts) <= 256


@pytest.mark.parametrize("world,cells", [(1, 500), (3, 500), (8, 4096), (2, (1 << 26))])
def test_frame_gather_is_a_stable_sort_by_grid_index(built, world, cells):
    """What a hook sees under sph_simulation::devices: the slabs' particles, stably sorted by
    grid_index with the slabs taken left to right (keys past the grid's bound in one last bucket) —
    the host threads' shared counting sort against numpy's stable argsort; the last case takes the
    one-thread path kept for very large grids."""
    from libclsph_b200 import PARTICLE, host
    rng = np.random.default_rng(world)
    n = 20000
    stage = np.zeros(n, PARTICLE)
    stage["position"][:, :3] = rng.normal(size=(n, 3)).astype(np.float32)
    keys = rng.integers(0, 600 if cells < (1 << 20) else 5000, n).astype(np.uint32)  # some beyond `cells`
    stage["grid_index"] = keys
    ids = rng.permutation(n).astype(np.uint32)
    cuts = np.sort(rng.integers(0, n + 1, world - 1)) if world > 1 else np.zeros(0, np.int64)
    start = np.concatenate([[0], cuts, [n]]).astype(np.uint32)  # slabs of uneven (possibly zero) size
    got, got_ids = host.gather_sort(stage, ids, start, cells)
    order = np.argsort(np.minimum(keys, cells), kind="stable")
    assert got.tobytes() == stage[order].tobytes()
    assert np.array_equal(got_ids, ids[order])


@pytest.mark.parametrize("world,threads", [(1, 1), (2, 2), (3, 3), (8, 8), (4, 1), (2, 5)])
def test_upload_deal_out_keeps_the_order_inside_a_slab(built, world, threads):
    """The upload side of sph_simulation::devices: every particle lands in the slab
    [cuts[r], cuts[r+1]) that holds its x (a particle exactly on a face belongs to the slab on its
    right), in the order it had in the caller's array, with its id beside it — the host threads'
    shared partition against a numpy stable argsort by slab."""
    from libclsph_b200 import PARTICLE, host
    rng = np.random.default_rng(10 * world + threads)
    n = 30011
    p = np.zeros(n, PARTICLE)
    p["position"][:, :3] = rng.normal(size=(n, 3)).astype(np.float32)
    p["velocity"][:, :3] = rng.normal(size=(n, 3)).astype(np.float32)
    ids = rng.permutation(n).astype(np.uint32)
    cuts = host.equal_count_cuts(p, world)
    if world > 1:
        p["position"][:17, 0] = cuts[1]  # on a face
    p["position"][17, 0], p["position"][18, 0] = np.float32(-1e30), np.float32(1e30)
    stage, stage_ids, start = host.deal_out(p, ids, cuts, threads)
    x = p["position"][:, 0]
    slab = np.clip(np.searchsorted(cuts[1:-1], x, side="right"), 0, world - 1)
    order = np.argsort(slab, kind="stable")
    assert np.array_equal(start, np.concatenate([[0], np.cumsum(np.bincount(slab, minlength=world))]).astype(np.uint32))
    assert stage.tobytes() == p[order].tobytes()
    assert np.array_equal(stage_ids, ids[order])
    for r in range(world):
        xs = stage["position"][start[r]:start[r + 1], 0]
        assert np.all((xs >= cuts[r]) & (xs < cuts[r + 1]))


def test_reference_caller_compiles_unchanged():
    """The reference's own example/particles.cpp (the boundary's caller) compiles
    UNCHANGED against libclsph_b200/host + include/ and links against the two product libraries
    (INTEGRATION.md section A).  Needs the reference tree: skipped where it is absent (GPU box)."""
    import subprocess
    from libclsph_b200 import build as b
    if not os.path.exists(os.path.join(b.REFERENCE, "example", "particles.cpp")):
        pytest.skip("no reference tree here")
    exe = b.build_reference_caller(force=True)
    assert exe and os.path.exists(exe)
    # too few arguments: the reference's usage text and its exit code (-1 -> 255); no engine is made
    r = subprocess.run([exe], capture_output=True, text=True, timeout=60)
    assert r.returncode == 255 and "Usage: ./sph <fluid_name>" in r.stdout


def test_geo_float_formatter_equals_printf_g():
    """The .geo writer formats floats itself (the reference's `stream << float` = printf "%g" was
    the end-to-end bottleneck of a saved frame): byte-identical to "%g" over 3 M values — random
    bit patterns (every exponent, denormals, infinities, NaNs), decimal round numbers, values one
    ulp around 6-digit rounding ties and around powers of ten."""
    from libclsph_b200 import host
    rng = np.random.default_rng(11)
    bits = rng.integers(0, 1 << 32, 2_000_000, dtype=np.uint64).astype(np.uint32)
    parts = [bits.view(np.float32)]
    dec = (rng.integers(1, 10 ** 6, 400_000) * 10.0 ** rng.integers(-12, 12, 400_000)).astype(np.float32)
    parts.append(dec)
    ties = ((rng.integers(10 ** 5, 10 ** 6, 200_000) + 0.5) * 10.0 ** rng.integers(-10, 6, 200_000)).astype(np.float32)
    for k in (-1, 0, 1):  # one ulp below, at, and above the float nearest to a tie
        parts.append((ties.view(np.int32) + k).view(np.float32))
    p10 = (10.0 ** np.arange(-37, 39)).astype(np.float32)
    for k in (-2, -1, 0, 1, 2):
        parts.append((p10.view(np.int32) + k).view(np.float32))
    parts.append(np.array([0.0, -0.0, 1e-45, -1e-45, 3.4028235e38, 1.17549435e-38, 999999.5, 999999.44, 0.0001,
                           0.00001, 100000.0, 1000000.0, 123456.5, 1234565.0, 0.5, 1.5, 2.5], np.float32))
    v = np.concatenate(parts)
    v = np.concatenate([v, -v])
    got = host.format_floats(v).split(b"\n")[:-1]
    assert len(got) == len(v)
    want = [("%g" % x).encode() for x in v.astype(np.float64)]
    # glibc prints the sign of a NaN ("-nan"); Python's % does not
    nan = np.isnan(v)
    for k in np.nonzero(nan)[0]:
        want[k] = b"-nan" if v.view(np.uint32)[k] >> 31 else b"nan"
    bad = [(float(v[k]), got[k], want[k]) for k in range(len(v)) if got[k] != want[k]]
    assert not bad, bad[:10]
