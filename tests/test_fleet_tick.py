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


def _restatement(oracle, px, py, dx, dy, vel, speed, dist, v_model, N, pid_state=None):
    idx, flag, (ox, oy, odx, ody, ov) = oracle.extract(px, py, dx, dy, vel, dist, layout="pb")
    if pid_state is None:                                # fresh slowDownCar per vehicle
        B0 = len(speed); pid_state = (np.zeros(B0, np.float32), np.zeros(B0, np.float32), np.ones(B0, np.uint8))
    oracle.crossing(px, py, vel, ov, pid_state)          # :445-473, the reference's config defaults
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


@pytest.mark.parametrize("layout", ["pb", "bp"])
def test_chunked_tick_equals_one_chunk(layout, monkeypatch):
    """A large tick is cut into 4 chunks whose uploads overlap the kernels of the chunk before (bmpc_fleet_tick_tobi);
    the cut must be invisible: every output bit-equal to the one-chunk tick, incl. ragged polylines, the device-side
    look-ups and the crossing PID state carried over two ticks."""
    from trajectory_controller_b200.engine import LOOKUP_DISTANCE, LOOKUP_VELOCITY, PID_SLOWDOWN, BatchedMPC
    B, P, N = 3001, 48, 4
    px, py, dx, dy, vel, speed, dist, _ = _stop_fleet(B, P, 31)
    npts = np.random.default_rng(4).integers(0, P + 1, B).astype(np.int32)
    arrs = [px, py, dx, dy, vel] if layout == "pb" else [np.ascontiguousarray(a.T) for a in (px, py, dx, dy, vel)]
    res = {}
    for mode, chunk_min in (("one", "1000000"), ("four", "1")):
        monkeypatch.setenv("BMPC_TICK_CHUNK_MIN", chunk_min)
        with BatchedMPC(2, 2, N, B) as eng:
            eng.set_lookup(LOOKUP_DISTANCE, [0.0, 1.0, 4.0], [0.4, 0.7, 1.6])
            eng.set_lookup(LOOKUP_VELOCITY, [0.0, 2.0, 5.0], [0.3, 2.0, 3.5])
            eng.set_pid(PID_SLOWDOWN, 1.0, 0.3, 0.05, 0.01)
            a = eng.fleet_tick_tobi(speed, None, *arrs, layout=layout, npoints=npts)      # look-ahead + velocity on the device
            b = eng.fleet_tick_tobi(speed, dist, *arrs, layout=layout, npoints=npts)      # second tick: PID state carried over
            res[mode] = (a, b, eng.get_pid(PID_SLOWDOWN, B))
    for t in (0, 1):
        for k, v in res["one"][t].items():
            w = res["four"][t][k]
            assert np.array_equal(v.view(np.uint8), w.view(np.uint8)), (layout, t, k)
    for v, w in zip(res["one"][2], res["four"][2]):
        assert np.array_equal(v.view(np.uint8), w.view(np.uint8))


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


def _stop_fleet(B, P, seed):
    """Polylines with 0, 1 or several stop points (velocity == 0) at random indices, ahead of and behind the car."""
    px, py, dx, dy, vel, speed, dist = _fleet(B, P, seed)
    rng = np.random.default_rng(seed + 1)
    vel[:] = 1.0
    nstop = rng.integers(0, 4, B)
    for b in range(B):
        vel[rng.integers(0, P, nstop[b]), b] = 0.0
    px = (px - np.float32(0.8) * rng.random(B).astype(np.float32)[None, :]).astype(np.float32)   # some points behind the car
    return px, py, dx, dy, vel, speed, dist, nstop


def test_crossing_clamp_and_pid_state_over_ticks(oracle):
    """(f)3: the stop-at-crossing clamp (src/...:445-473) with one slowDownCar PID per vehicle resident on the device.
    Non-trivial gains make the state matter: three ticks on moving polylines, ragged point counts, then reset()."""
    from trajectory_controller_b200.engine import PID_SLOWDOWN, BatchedMPC, CrossingConfig
    B, P, N = 3000, 40, 4
    gains = (0.8, 0.5, 0.02, 0.01); cfgv = (0.9, 0.3, 1.2, 0.07)
    px, py, dx, dy, vel, speed, dist, nstop = _stop_fleet(B, P, 33)
    npts = np.random.default_rng(4).integers(0, P + 1, B).astype(np.int32); npts[::3] = P
    st = (np.zeros(B, np.float32), np.zeros(B, np.float32), np.ones(B, np.uint8))
    with BatchedMPC(2, 2, N, B) as eng:
        eng.set_pid(PID_SLOWDOWN, *gains)
        eng.set_crossing(CrossingConfig(*cfgv))
        seen = set()
        for tick in range(4):
            if tick == 3:                                 # PID::reset() for the whole fleet
                eng.reset_pid(PID_SLOWDOWN)
                st = (np.zeros(B, np.float32), np.zeros(B, np.float32), np.ones(B, np.uint8))
            pxt = (px - np.float32(0.12 * tick)).astype(np.float32)      # the car approaches the stop points
            got = eng.fleet_tick_tobi(speed, dist, pxt, py, dx, dy, vel, layout="pb", npoints=npts)
            idx, flag, (ox, oy, odx, ody, ov) = oracle.extract(pxt, py, dx, dy, vel, dist, npoints=npts, layout="pb")
            raw = ov.copy()
            oracle.crossing(pxt, py, vel, ov, st, gains=gains, cfg=cfgv, npoints=npts)
            assert np.array_equal(got["target_speed"].view(np.uint32), ov.view(np.uint32)), tick
            gi, gp, gf = eng.get_pid(PID_SLOWDOWN, B)
            assert np.array_equal(gi.view(np.uint32), st[0].view(np.uint32)) and np.array_equal(gp.view(np.uint32), st[1].view(np.uint32))
            assert np.array_equal(gf, st[2]), tick
            assert np.array_equal(got["state"], (ov != 0).astype(np.uint8))
            cross = ov < np.float32(0.5)
            assert np.all(got["steering_front"][cross] == 0) and np.all(got["steering_rear"][cross] == 0)
            seen |= {"clamped" if (ov != raw).any() else "", "zero" if ((ov == 0) & (raw != 0)).any() else "",
                     "partial" if ((ov > 0) & (ov < np.float32(cfgv[0])) & (ov != raw)).any() else ""}
        assert {"clamped", "zero", "partial"} <= seen
        eng.set_crossing(None)                            # clamp off: the polyline's own velocity comes back
        got = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb", npoints=npts)
        _, _, (_, _, _, _, ov) = oracle.extract(px, py, dx, dy, vel, dist, npoints=npts, layout="pb")
        assert np.array_equal(got["target_speed"].view(np.uint32), ov.view(np.uint32))


def test_lookup_tables_on_device(oracle):
    """LookupTable::linearSearch (src/...:66, :323) evaluated on the device: equals the restatement bit for bit, and a
    tick with dist=None / device velocity table equals the tick fed with the host-evaluated values."""
    from trajectory_controller_b200.engine import LOOKUP_DISTANCE, LOOKUP_VELOCITY, BatchedMPC, BmpcError
    B, P, N = 2000, 48, 4
    px, py, dx, dy, vel, speed, _ = _fleet(B, P, 17)
    dvx, dvy = [0.0, 0.5, 1.0, 2.0, 3.0], [0.45, 0.5, 0.62, 0.9, 1.3]
    vvx, vvy = [0.1, 1.0, 2.5], [0.3, 1.0, 2.0]
    with BatchedMPC(2, 2, N, B) as eng:
        with pytest.raises(BmpcError):
            eng.fleet_tick_tobi(speed, None, px, py, dx, dy, vel, layout="pb")      # no table, no FOH
        with pytest.raises(BmpcError):
            eng.set_lookup(LOOKUP_DISTANCE, [0.0, 0.0], [1.0, 2.0])                # not ascending
        eng.set_lookup(LOOKUP_DISTANCE, dvx, dvy)
        xs = np.concatenate([speed, np.float32([-1, 0, 0.5, 3, 7, 1.0])])
        assert np.array_equal(eng.lookup(LOOKUP_DISTANCE, xs).view(np.uint32), oracle.lookup(dvx, dvy, xs).view(np.uint32))
        dist = oracle.lookup(dvx, dvy, speed)
        a = eng.fleet_tick_tobi(speed, None, px, py, dx, dy, vel, layout="pb")
        b = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb")
        for k in a:
            assert np.array_equal(a[k], b[k]), k
        # FOH drive mode, src/...:70-73 in float
        eng.set_drive_mode(True, 0.6, 0.1)
        a = eng.fleet_tick_tobi(speed, None, px, py, dx, dy, vel, layout="pb")
        b = eng.fleet_tick_tobi(speed, (np.float32(0.6) + speed * np.float32(0.1)).astype(np.float32), px, py, dx, dy, vel, layout="pb")
        for k in a:
            assert np.array_equal(a[k], b[k]), k
        eng.set_drive_mode(False)
        # velocity table: v (after the |v| < 0.1 rule, a double) -> float -> table -> double  (:323)
        eng.set_lookup(LOOKUP_VELOCITY, vvx, vvy)
        v = speed.astype(np.float64); v = np.where(np.abs(v) < 0.1, 0.1, v)
        vm = oracle.lookup(vvx, vvy, v.astype(np.float32)).astype(np.float64)
        a = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb")
        b = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb", v_model=vm)
        for k in a:
            assert np.array_equal(a[k], b[k]), k
        eng.set_lookup(LOOKUP_VELOCITY, [], [])
        c = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel, layout="pb")
        assert not np.array_equal(c["steering_front"], a["steering_front"])


def test_fleet_tick_pid_branch(oracle):
    """The non-MPC branch (src/...:214-225): steering = pidFront(position.y), pidRear(directory.angle()), with the
    per-vehicle PID states resident across ticks.  steering_front is exact; steering_rear goes through atan2f (CUDA's,
    <= 2 ulp against libm): exact wherever the angle agrees, otherwise within the PID gain times the angle difference."""
    from trajectory_controller_b200.engine import PID_FRONT, PID_REAR, BatchedMPC
    B, P = 2500, 40
    px, py, dx, dy, vel, speed, dist, _ = _stop_fleet(B, P, 51)
    gf, gr = (1.2, 0.7, 0.03, 0.01), (0.9, 1.1, 0.01, 0.01)
    sf_st = (np.zeros(B, np.float32), np.zeros(B, np.float32), np.ones(B, np.uint8))
    sr_st = (np.zeros(B, np.float32), np.zeros(B, np.float32), np.ones(B, np.uint8))
    sd_st = (np.zeros(B, np.float32), np.zeros(B, np.float32), np.ones(B, np.uint8))
    with BatchedMPC(4, 1, 10, B) as eng:                  # any handle: no MPC shape involved
        eng.set_pid(PID_FRONT, *gf); eng.set_pid(PID_REAR, *gr)
        for tick in range(3):
            pyt = (py + np.float32(0.01 * tick)).astype(np.float32)
            got = eng.fleet_tick_pid(speed, dist, px, pyt, dx, dy, vel, layout="pb")
            idx, flag, (ox, oy, odx, ody, ov) = oracle.extract(px, pyt, dx, dy, vel, dist, layout="pb")
            oracle.crossing(px, pyt, vel, ov, sd_st)
            sf, sr, td = oracle.pid_branch(ox, oy, odx, ody, gf, gr, sf_st, sr_st)
            cross = ov < np.float32(0.5)
            sf = np.where(cross, 0.0, sf); sr = np.where(cross, 0.0, sr)
            assert np.array_equal(got["idx"], idx) and np.array_equal(got["flag"], flag)
            assert np.array_equal(got["target_speed"].view(np.uint32), ov.view(np.uint32))
            assert np.array_equal(got["target_distance"].view(np.uint32), td.view(np.uint32))
            assert np.array_equal(got["steering_front"], sf), tick
            assert np.max(np.abs(got["steering_rear"] - sr)) <= 1e-4, np.max(np.abs(got["steering_rear"] - sr))
            assert (got["steering_rear"] == sr).mean() > 0.5
            gi, gp, gfst = eng.get_pid(PID_FRONT, B)
            assert np.array_equal(gi.view(np.uint32), sf_st[0].view(np.uint32)) and np.array_equal(gp.view(np.uint32), sf_st[1].view(np.uint32))
            ri, rp, _ = eng.get_pid(PID_REAR, B)          # carry the device's rear state into the restatement so that an
            sr_st[0][:] = ri; sr_st[1][:] = rp            # atan2f ulp does not accumulate through the integral term
