"""GPU: the C++11 host classes end to end — sph_simulation::simulate() (libclsph_b200/host/) drives
the engine through load_settings, scene::load, the lattice, the frame / sub-step loop, the save
thread and the Houdini writer — against the same steps made through the C ABI directly, and the
command-line driver built from host/example/particles.cpp."""
import os
import subprocess

import numpy as np
import pytest

from common import DEFAULT_SIM, FLUIDS, ROOT, box_mesh, make_case, new_engine, write_json, write_obj

pytestmark = pytest.mark.gpu


def test_simulate_matches_direct_stepping(tmp_path):
    from libclsph_b200 import host
    case = make_case(tmp_path, "water", 8000)
    steps = 7
    frames = tmp_path / "out"
    os.makedirs(frames / "frames")
    r = host.simulate(case["fluid_json"], case["sim_json"], case["obj"], max_steps=steps,
                      frames_prefix=str(frames) + "/", n=8000)
    assert r["steps"] == steps
    assert r["frames"] == 2  # the initial state + the (cut short) first frame
    # the same 7 sub-steps through the C ABI, with the frame loop's dt bookkeeping
    with new_engine(case) as e:
        e.upload(case["particles"])
        timeperframe = np.float32(1.0) / case["params"]["target_fps"][0]
        dt = np.float32(timeperframe * case["params"]["simulation_scale"][0])
        timeleft = timeperframe
        for _ in range(steps):
            dt = np.float32(e.step(float(dt)))
            timeleft = np.float32(timeleft - dt)
            if timeleft < dt:
                dt = timeleft
        want = e.download()
        wantp = e.parameters()
    got = r["state"]
    for f in ("position", "velocity", "intermediate_velocity", "acceleration", "density", "pressure", "grid_index"):
        assert np.array_equal(got[f].view(np.uint32), want[f].view(np.uint32)), f
    assert r["params"].tobytes() == wantp.tobytes()
    # two .geo frames, the second holding the final state (byte-identical to the writer's text)
    names = sorted(os.listdir(frames / "frames"))
    assert names == ["frame0000001.geo", "frame0000002.geo"]
    text = open(frames / "frames" / names[1], "rb").read()
    assert text == host.geo_text(got, float(case["params"]["particle_mass"][0]))


def test_command_line_driver(tmp_path):
    """clsphparticles <fluid> <sim> <scene> <prefix> -y in a directory laid out like the
    reference's (fluid_properties/, simulation_properties/, scenes/), 2 frames of 512 particles."""
    exe = os.path.join(ROOT, "libclsph_b200", "clsphparticles")
    assert os.path.exists(exe)
    for d in ("fluid_properties", "simulation_properties", "scenes", "frames"):
        os.makedirs(tmp_path / d)
    write_json(str(tmp_path / "fluid_properties" / "water.json"), FLUIDS["water"])
    write_json(str(tmp_path / "simulation_properties" / "tiny.json"),
               dict(DEFAULT_SIM, particles_count=512, simulation_time=0.03, serialize=True))
    v, f = box_mesh(-1, 1, -2, 0, -1, 1)
    write_obj(str(tmp_path / "scenes" / "box.obj"), [(v, f)])
    r = subprocess.run([exe, "water", "tiny", "box.obj", "", "-y"], cwd=tmp_path, capture_output=True, text=True,
                       timeout=300)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-2000:]
    assert "Simulating frame 2" in r.stdout and "Duration :" in r.stdout and "totalgpoints" in r.stdout
    geo = sorted(os.listdir(tmp_path / "frames"))
    assert geo[0] == "frame0000001.geo" and len(geo) == 3  # initial + 2 frames of 1/60 s
    assert os.path.getsize(tmp_path / "last_frame.bin") == 512 * 80
    head = open(tmp_path / "frames" / geo[-1]).read(40)
    assert head.startswith("PGEOMETRY V5\nNPoints 512 NPrims 1")
