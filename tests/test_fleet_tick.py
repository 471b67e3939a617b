"""bmpc_fleet_tick_tobi: the tobiMPC branch of TrajectoryPointController::cycle()
(/root/reference/src/trajectory_point_follower.cpp:64-126, 276-283, 301-389) for a fleet, against a CPU restatement
composed from the pinned pieces: oracle point extraction, libm atan2, the reference's model formulas in numpy doubles
(each a single IEEE operation), and the oracle's cold dlib::mpc<2,2,4>.   [-m gpu]

Bars: polyline index / flag and the trajectory point's floats bit-exact; y_soll bit-exact; phi_soll within 2 ulp of
libm's atan2 (CUDA's double atan2 is specified to 2 ulp); steering within 1e-9 rad; target_speed / target_distance /
state / the crossing rule exact."""
import numpy as np
import pytest

from trajectory_controller_b200 import workloads as W

pytestmark = pytest.mark.gpu


def _fleet(B, P, seed):
    """Vehicle-frame polylines [P][B]: gentle random curves ahead of the car, per-point speeds, some stop points."""
    rng = np.random.default_rng(seed)
    s = np.arange(P, dtype=np.float64)[:, None] * 0.1
    curv = rng.uniform(-0.8, 0.8, B)[None, :]; off = rng.uniform(-0.15, 0.15, B)[None, :]; head = rng.uniform(-0.3, 0.3, B)[None, :]
    ang = head + curv * s
    x = np.cumsum(0.1 * np.cos(ang), axis=0); y = off + np.cumsum(0.1 * np.sin(ang), axis=0)
    vel = np.full((P, B), 1.0); vel[:, rng.random(B) < 0.2] = 0.3; vel[:, rng.random(B) < 0.1] = 0.0
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    speed = rng.uniform(-0.05, 2.5, B).astype(np.float32)           # includes |v| < 0.1
    dist = (np.float32(0.6) + speed * np.float32(0.1)).astype(np.float32)   # :71, float arithmetic
    return f(x), f(y), f(np.cos(ang)), f(np.sin(ang)), f(vel), speed, dist


def _restatement(oracle, px, py, dx, dy, vel, speed, dist, v_model, N):
    idx, flag, (ox, oy, odx, ody, ov) = oracle.extract(px, py, dx, dy, vel, dist, layout="pb")
    v = speed.astype(np.float64); v = np.where(np.abs(v) < 0.1, 0.1, v)
    phi = np.arctan2(ody.astype(np.float64), odx.astype(np.float64)); ys = oy.astype(np.float64)
    if v_model is not None:
        v = v_model
    B = len(v); T, l = 0.1, 0.21
    Tv = T * v; Tvl = Tv / l
    A = np.zeros((B, 2, 2)); A[:, 0, 0] = 1; A[:, 0, 1] = Tv; A[:, 1, 1] = 1
    Bm = np.zeros((B, 2, 2)); Bm[:, 0, 1] = Tv; Bm[:, 1, 0] = Tvl; Bm[:, 1, 1] = -Tvl
    am = 22 * np.pi / 180
    models = dict(A=A, B=Bm, C=np.zeros((B, 2)), Q=np.tile([20.0, 7.0], (B, 1)), R=np.tile([0.0005, 10.0], (B, 1)),
                  lower=np.full((B, 2), -am), upper=np.full((B, 2), am))
    tgt = np.stack([ys, phi], axis=1)
    u0, _, it, _ = oracle.batch_cold(models, N, np.zeros((B, 2)), target=tgt, per_instance=True, nthreads=4, want_controls=False)
    cross = ov < np.float32(0.5)
    sf = np.where(cross, 0.0, u0[:, 0]); sr = np.where(cross, 0.0, u0[:, 1])
    td = np.sqrt((ox * ox + oy * oy).astype(np.float32)).astype(np.float32)
    return dict(idx=idx, flag=flag, steering_front=sf, steering_rear=sr, target_speed=ov, target_distance=td,
                state=(ov != 0).astype(np.uint8), target=tgt, iters=it)


@pytest.mark.parametrize("B,use_lookup", [(5000, False), (777, True)])
def test_fleet_tick_matches_restatement(oracle, B, use_lookup):
    from trajectory_controller_b200.engine import BatchedMPC
    P, N = 48, 4
    px, py, dx, dy, vel, speed, dist = _fleet(B, P, 9)
    v_model = None
    if use_lookup:      # a caller-side velocity lookup (:323): piecewise-linear table evaluated on the host
        v_model = np.interp(np.abs(speed.astype(np.float64)), [0.0, 1.0, 3.0], [0.3, 1.0, 2.2])
    want = _restatement(oracle, px, py, dx, dy, vel, speed, dist, v_model, N)
    with BatchedMPC(2, 2, N, B) as eng:
        got = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb", v_model=v_model)
        pattern = eng.pattern_name()
        got2 = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb", v_model=v_model)   # fresh controllers each tick
    assert pattern == "tobi"
    for k in ("idx", "flag", "target_speed", "state"):
        assert np.array_equal(got[k], want[k]), k
    assert np.array_equal(got["target_distance"].view(np.uint32), want["target_distance"].view(np.uint32))
    assert np.array_equal(got["target"][:, 0].view(np.uint64), want["target"][:, 0].view(np.uint64))       # y_soll
    dphi = np.abs(got["target"][:, 1] - want["target"][:, 1])
    assert np.all(dphi <= 2 * np.spacing(np.abs(want["target"][:, 1]))), dphi.max()
    for k in ("steering_front", "steering_rear"):
        assert np.max(np.abs(got[k] - want[k])) <= 1e-9, (k, np.max(np.abs(got[k] - want[k])))
        assert np.array_equal(got[k].view(np.uint64), got2[k].view(np.uint64))
    same_phi = got["target"][:, 1] == want["target"][:, 1]
    assert same_phi.mean() > 0.5
    # where atan2 agrees bit for bit everything downstream is exact as well
    cross = want["target_speed"] < 0.5
    assert np.array_equal(got["iters"][same_phi], want["iters"][same_phi])
    for k in ("steering_front", "steering_rear"):
        assert np.array_equal(got[k][same_phi], want[k][same_phi])
    assert np.all(got["steering_front"][cross] == 0) and np.all(got["steering_rear"][cross] == 0) and cross.any()
    assert np.all(np.abs(got["steering_front"]) <= 22 * np.pi / 180 + 1e-15)


def test_fleet_tick_rejects_wrong_shape():
    from trajectory_controller_b200.engine import BatchedMPC, BmpcError
    z = np.zeros((4, 3), dtype=np.float32)
    with BatchedMPC(4, 1, 10, 8) as eng:
        with pytest.raises(BmpcError):
            eng.fleet_tick_tobi(np.ones(3, dtype=np.float32), np.ones(3, dtype=np.float32), z, z, z, z, z)


def test_fleet_tick_ragged_and_empty_polylines(oracle):
    """npoints per vehicle: 0 (empty TRAJECTORY -> the default point (d,0),(1,0), velocity 0: src/...:394-407 -> target 0,
    IDLE, zero steering), 1 (only the fallback point) and short polylines that end before the look-ahead distance."""
    from trajectory_controller_b200.engine import BatchedMPC
    B, P, N = 600, 48, 4
    px, py, dx, dy, vel, speed, dist = _fleet(B, P, 21)
    npts = np.random.default_rng(2).integers(0, P + 1, B).astype(np.int32)
    npts[:20] = 0; npts[20:40] = 1; npts[40:60] = 3
    idx, flag, (ox, oy, odx, ody, ov) = oracle.extract(px, py, dx, dy, vel, dist, npoints=npts, layout="pb")
    with BatchedMPC(2, 2, N, B) as eng:
        got = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb", npoints=npts)
    assert np.array_equal(got["idx"], idx) and np.array_equal(got["flag"], flag)
    assert np.array_equal(got["target_speed"].view(np.uint32), ov.view(np.uint32))
    empty = npts == 0
    assert np.all(got["state"][empty] == 0) and np.all(got["steering_front"][empty] == 0) and np.all(got["steering_rear"][empty] == 0)
    assert np.all(got["target"][empty] == 0) and np.all(got["iters"][empty] == 0)      # already converged: 0 iterations
    assert np.array_equal(got["target_distance"][empty].view(np.uint32), dist[empty].view(np.uint32))   # |(d, 0)| = d
    assert np.array_equal(got["target"][:, 0].view(np.uint64), oy.astype(np.float64).view(np.uint64))
