"""Closed loop on the device (BASELINE.json configs[4]): extraction -> target -> warm solve ->
plant, all in kernels, against the CPU restatement oracle/closed_loop_oracle.c (persistent orc_mpc per vehicle)."""
import math

import numpy as np
import pytest

from _oracle import same_values
from trajectory_controller_b200 import workloads as W


def test_harness_atan_is_an_arctangent(oracle):
    """The harness arctangent (include/bmpc.h) is accurate to a few ulp on the range the loop uses."""
    f = oracle.lib.orc_harness_atan
    import ctypes as C
    f.restype = C.c_double; f.argtypes = [C.c_double]
    for t in np.linspace(-1, 1, 401):                      # |dir.y/dir.x| <= 1: headings within 45 degrees of the path
        assert abs(f(float(t)) - math.atan(t)) < 5e-16
    for t in np.linspace(-3, 3, 601):                      # the truncated series degrades gracefully outside
        assert abs(f(float(t)) - math.atan(t)) < 1e-12


def test_oracle_closed_loop_tracks_the_path(oracle):
    """Sanity of the harness itself on the CPU: vehicles converge onto the sine path."""
    path = W.sine_world_path()
    pose0 = W.closed_loop_poses(64, 45, path)
    r = oracle.closed_loop(W.bicycle4_model(), 30, path, pose0, ticks=150)
    # lateral distance to the path after 15 m: nearest path point
    d = np.min(np.hypot(r["pose"][:, 0, None] - path[0][None, :], r["pose"][:, 1, None] - path[1][None, :]), axis=1)
    assert np.all(d < 0.12), d.max()
    assert np.all(r["iters_max"] <= 10000) and np.all(r["iters_total"] > 0)


@pytest.mark.gpu
@pytest.mark.parametrize("N,count,ticks", [(30, 768, 40), (10, 1000, 120)])
def test_closed_loop_matches_oracle(oracle, N, count, ticks):
    from trajectory_controller_b200.engine import BatchedMPC
    path = W.sine_world_path()
    pose0 = W.closed_loop_poses(count, 45, path)
    bike = W.bicycle4_model()
    want = oracle.closed_loop(bike, N, path, pose0, ticks=ticks)
    with BatchedMPC(4, 1, N, count) as eng:
        eng.set_model(bike)
        eng.cl_init(path, pose0)
        eng.cl_run(ticks // 2)
        eng.cl_run(ticks - ticks // 2)                     # the loop can be resumed
        got = eng.cl_get()
    for k in ("pose", "act", "u0", "target"):
        assert same_values(got[k], want[k]), k
    assert np.array_equal(got["idx"], want["idx"])
    assert np.array_equal(got["iters_total"], want["iters_total"]) and np.array_equal(got["iters_max"], want["iters_max"])


@pytest.mark.gpu
def test_closed_loop_full_size_properties():
    """configs[4] at full width (262,144 vehicles; 60 ticks here): every vehicle ends near the path, results do not
    depend on where in the batch a vehicle sits, steering stays inside the box."""
    from trajectory_controller_b200.engine import BatchedMPC
    count, ticks = 262144, 60
    path = W.sine_world_path()
    pose0 = W.closed_loop_poses(count, 45, path)
    with BatchedMPC(4, 1, 30, count) as eng:
        eng.set_model(W.bicycle4_model())
        eng.cl_init(path, pose0)
        eng.cl_run(ticks)
        a = eng.cl_get()
        perm = np.random.default_rng(0).permutation(count)
        eng.cl_init(path, pose0[perm])
        eng.cl_run(ticks)
        b = eng.cl_get()
    assert np.array_equal(a["pose"][perm].view(np.uint64), b["pose"].view(np.uint64))
    assert np.array_equal(a["iters_total"][perm], b["iters_total"])
    assert np.all(np.abs(a["u0"]) <= W.ALPHA_MAX)
    sub = a["pose"][:: 64]
    d = np.min(np.hypot(sub[:, 0, None] - path[0][None, :], sub[:, 1, None] - path[1][None, :]), axis=1)
    assert np.all(d < 0.2), d.max()


def test_oracle_horizon_targets_track_better(oracle):
    """A target per horizon step follows the sine path more tightly than one look-ahead target."""
    path = W.sine_world_path()
    pose0 = W.closed_loop_poses(48, 45, path)
    bike = W.bicycle4_model()
    def err(r):
        return np.min(np.hypot(r["pose"][:, 0, None] - path[0][None, :], r["pose"][:, 1, None] - path[1][None, :]), axis=1)
    a = oracle.closed_loop(bike, 30, path, pose0, ticks=200)
    b = oracle.closed_loop(bike, 30, path, pose0, ticks=200, horizon=(0.1, 0.1))
    assert err(b).mean() < err(a).mean()


@pytest.mark.gpu
@pytest.mark.parametrize("N,count,ticks,hz", [(30, 700, 36, (0.1, 0.1)), (10, 900, 90, (0.7, 0.05)), (30, 300, 20, (0.6, 0.0))])
def test_closed_loop_horizon_targets_match_oracle(oracle, N, count, ticks, hz):
    """The loop with per-step targets sampled on the device (k_cl_prepare_horizon -> target ring -> warm solve) against
    the CPU restatement with set_target(val, t) per step (orc_closed_loop_h): bit for bit."""
    from trajectory_controller_b200.engine import BatchedMPC
    path = W.sine_world_path()
    pose0 = W.closed_loop_poses(count, 46, path)
    bike = W.bicycle4_model()
    want = oracle.closed_loop(bike, N, path, pose0, ticks=ticks, horizon=hz)
    with BatchedMPC(4, 1, N, count) as eng:
        eng.set_model(bike)
        eng.cl_init(path, pose0)
        eng.cl_set_horizon_targets(True, *hz)
        eng.cl_run(ticks // 3)
        eng.cl_run(ticks - ticks // 3)
        got = eng.cl_get()
        # switching back gives the reference's single-target loop again
        eng.cl_init(path, pose0)
        eng.cl_run(5)
        plain = eng.cl_get()
    for k in ("pose", "act", "u0", "target"):
        assert same_values(got[k], want[k]), k
    assert np.array_equal(got["idx"], want["idx"])
    assert np.array_equal(got["iters_total"], want["iters_total"]) and np.array_equal(got["iters_max"], want["iters_max"])
    ref5 = oracle.closed_loop(bike, N, path, pose0, ticks=5)
    assert same_values(plain["pose"], ref5["pose"])
