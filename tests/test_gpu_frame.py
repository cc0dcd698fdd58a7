"""GPU tests of the device-driven sub-step loop (sphb_run_frame): a frame's
steps enqueued without host round trips must give the BYTES that calling sphb_step in the
reference's loop (libclsph/sph_simulation.cpp:384-409) gives — same arithmetic, only where the
host half of a step is evaluated differs."""
import os
import tempfile

import numpy as np
import pytest

from common import GOLDEN, make_case, new_engine

pytestmark = pytest.mark.gpu

FRAME = np.float32(1.0) / np.float32(60.0)


def _host_loop(e, dt, frame_time, max_steps=0):
    """the reference's loop over sphb_step, in fp32 like the C++ caller"""
    dt = np.float32(dt)
    timeleft = np.float32(frame_time)
    log, attempts = [], []
    while timeleft > 0.0:
        if max_steps and len(log) >= max_steps:
            break
        dt = np.float32(e.step(float(dt)))
        attempts.append(e.debug()["attempts"])
        timeleft = np.float32(timeleft - dt)
        if timeleft < dt:
            dt = timeleft
        log.append(timeleft)
    return float(dt), np.array(log, np.float32), attempts


def _seeded_state(case, seed, speed):
    ps = case["particles"].copy()
    rng = np.random.default_rng(seed)
    v = rng.uniform(-speed, speed, (len(ps), 3)).astype(np.float32)
    ps["velocity"][:, :3] = v
    ps["intermediate_velocity"][:, :3] = v
    return ps


def _compare(case, particles, frames, options, expect_queued=True, max_steps=0):
    dt0 = float(np.float32(FRAME) * np.float32(case["params"]["simulation_scale"][0]))
    with new_engine(case, 0) as a, new_engine(case, 0) as b:
        for name, value in options.items():
            a.set_option(name, value)
            b.set_option(name, value)
        a.upload(particles)
        b.upload(particles)
        dta = dtb = dt0
        all_attempts = []
        for f in range(frames):
            dta, log_a, attempts = _host_loop(a, dta, FRAME, max_steps)
            all_attempts += attempts
            dtb, log_b, steps = b.run_frame(dtb, float(FRAME), max_steps)
            assert steps == len(log_a), f"frame {f}: {steps} steps vs {len(log_a)}"
            assert np.array_equal(log_a.view(np.uint32), log_b.view(np.uint32)), f"frame {f}: timeleft log"
            assert np.float32(dta).tobytes() == np.float32(dtb).tobytes(), f"frame {f}: dt"
            assert a.parameters().tobytes() == b.parameters().tobytes(), f"frame {f}: parameters"
            assert a.download().tobytes() == b.download().tobytes(), f"frame {f}: state"
            da, db = a.debug(), b.debug()
            for k in ("keys", "perm", "cell_table"):
                assert np.array_equal(da[k], db[k]), f"frame {f}: {k}"
        queued = b.queued_steps()
        if expect_queued:
            assert queued > 0
        return all_attempts, queued


@pytest.mark.parametrize("graph", [0, 1])
def test_run_frame_matches_stepping_dambreak(graph):
    # the collapsing block: dt shrinks from the first steps on, so most steps retry the advection
    with tempfile.TemporaryDirectory() as tmp:
        case = make_case(tmp, "water", 4096)
        attempts, queued = _compare(case, case["particles"], 3, {"graph": graph})
        assert max(attempts) >= 2          # the retry ran on the device
        # all but the first step and the hand-overs (table growth while the block spreads, a third attempt)
        assert queued >= 0.8 * len(attempts)


def test_run_frame_host_takes_over_attempts():
    # one enqueued attempt per step: every retry is finished by the host
    with tempfile.TemporaryDirectory() as tmp:
        case = make_case(tmp, "water", 4096)
        attempts, _ = _compare(case, case["particles"], 2, {"queued_attempts": 1})
        assert max(attempts) >= 2


@pytest.mark.parametrize("graph", [0, 1])
def test_run_frame_host_grows_tables(graph):
    # an expanding cloud with tables allocated exactly: every grid growth goes through the host
    with tempfile.TemporaryDirectory() as tmp:
        case = make_case(tmp, "water", 2197)
        ps = _seeded_state(case, 5, 4.0)
        with new_engine(case, 0) as probe:
            probe.upload(ps)
            probe.step(1e-3)
            g0 = probe.parameters()[["grid_size_x", "grid_size_y", "grid_size_z"]].copy()
            for _ in range(25):
                probe.step(1.0 / 600)
            g1 = probe.parameters()[["grid_size_x", "grid_size_y", "grid_size_z"]]
            assert g0.tobytes() != g1.tobytes(), "the cloud must outgrow its first grid for this test"
        _compare(case, ps, 3, {"table_headroom": 0, "graph": graph})


def test_run_frame_max_steps_and_options():
    with tempfile.TemporaryDirectory() as tmp:
        case = make_case(tmp, "mucus", 2744)
        _compare(case, case["particles"], 2, {}, max_steps=4)
        # options that fall back to per-step stepping still give the same answers
        _compare(case, case["particles"], 1, {"device_loop": 0}, expect_queued=False)
        _compare(case, case["particles"], 1, {"sort": 1}, expect_queued=False)


def test_run_frame_mucus_monkeybox_fixture_scene():
    # 978 faces: the triangle bin inside a captured step
    from libclsph_b200 import Engine
    z = np.load(os.path.join(GOLDEN, "steps_monkeybox_mucus.npz"))
    sc = np.load(os.path.join(GOLDEN, "scene_monkeybox_mucus.npz"))
    case = dict(params=z["params0"], terms=z["terms"], particles=z["initial"],
                scene=dict(bbs=sc["bbs"], transforms=sc["transforms"], rvertices=sc["rvertices"],
                           total_gridpoints=int(sc["total_gridpoints"])))
    for graph in (0, 1):
        _compare(case, case["particles"], 2, {"graph": graph})


def test_two_engines_in_run_frame_at_once(tmp_path):
    """Two engines on ONE device inside sphb_run_frame at the same time, from two threads: the
    device-driven steps share per-process constants, so one engine at a time holds the device's
    lease and the other steps from the host for that batch — both must give exactly what each
    gives alone (different particle counts, so a leaked grid would be visible at once)."""
    import threading
    for d in ("a", "b"):
        os.makedirs(tmp_path / d)
    cases = [make_case(tmp_path / "a", "water", 32768), make_case(tmp_path / "b", "water", 46656)]
    alone = []
    for c in cases:
        with new_engine(c) as e:
            e.set_option("device_loop", 1)
            e.upload(c["particles"])
            e.step(1.0 / 600.0)
            for _ in range(4):
                e.run_frame(1.0 / 600.0, 1.0, 3)
            alone.append(e.download())
    for attempt in range(3):
        engines = [new_engine(c) for c in cases]
        got = [None, None]

        def work(k):
            e = engines[k]
            e.set_option("device_loop", 1)
            e.upload(cases[k]["particles"])
            e.step(1.0 / 600.0)
            for _ in range(4):  # four calls of three steps: the lease changes hands between them
                e.run_frame(1.0 / 600.0, 1.0, 3)
            got[k] = e.download()

        threads = [threading.Thread(target=work, args=(k,)) for k in range(2)]
        for t in threads:
            t.start()
        for t in threads:
            t.join()
        for e in engines:
            e.close()
        for k in range(2):
            assert got[k].tobytes() == alone[k].tobytes(), (attempt, k)
