"""Parity of the CUDA path, called through the C ABI, against the golden vectors of the real dlib::mpc and
against the pinned oracle.   [-m gpu]

Bar (EXACT mode, FP64): IEEE value equality of u0 and of the whole control sequence (the sign of a zero is the only
freedom: the kernel does not multiply by diagm()'s literal zeros) and IDENTICAL iteration counts.
"""
import numpy as np
import pytest

from _oracle import same_values
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import EXACT, FP64, BatchedMPC

pytestmark = pytest.mark.gpu


def _cold(model, N, x0, target=None, eps=0.01, max_iter=10000, per_instance=False):
    m = model
    S = (m["B"] if isinstance(m, dict) else m.B).shape[-2]
    I = (m["B"] if isinstance(m, dict) else m.B).shape[-1]
    with BatchedMPC(S, I, N, len(x0), FP64, EXACT) as eng:
        eng.set_model(m, per_instance=per_instance)
        eng.set_options(eps, max_iter)
        if target is not None:
            eng.set_target(target)
        u0, it, st = eng.solve(x0, keep_controls=True)
        return u0, eng.get_controls(len(x0)), it, st, eng.derived(0)


# ------------------------------------------------------------------------------------------ golden (real dlib)
@pytest.mark.parametrize("tag,eps", [("kat_tight", 1e-8), ("kat_default", 0.01)])
def test_kat_closed_loop_matches_dlib(golden, tag, eps):
    """dlib's own known-answer scenario (dlib/test/mpc.cpp:269-316) through a persistent engine slot; the batch holds
    3 copies of the controller to check slots do not interact."""
    kat = W.kat_model()
    with BatchedMPC(2, 1, 30, 3) as eng:
        eng.set_model(kat)
        eng.set_options(eps, 10000)
        lam, qd = eng.derived()
        assert lam == golden[tag + "_lambda"][0] and same_values(qd, golden[tag + "_qdiag"])
        for k in range(30):
            x = np.tile(golden[tag + "_x"][k], (3, 1))
            u0, it, st = eng.solve(x, warm=True)
            for b in range(3):
                assert same_values(u0[b], golden[tag + "_u0"][k]), (k, b)
                assert it[b] == golden[tag + "_iters"][k], (k, b)
    if tag == "kat_default":
        assert u0[0, 0] == -0.10000002376964306


def test_tobi_module_configuration(golden):
    """The reference module's own controller: mpc<2,2,4>, fresh per cycle (src/trajectory_point_follower.cpp:366-380)."""
    u0, _, _, _, _ = _cold(W.tobi_model(1.0), 4, np.zeros((1, 2)), target=np.array([[0.05, 0.1]]))
    assert tuple(u0[0]) == (0.27357241694602841, 0.023830246920932288)
    models = [W.tobi_model(float(v)) for v in golden["tobi_rand_v"]]
    pm = {k: np.stack([getattr(m, k) for m in models]) for k in ("A", "B", "C", "Q", "R", "lower", "upper")}
    u0, ctr, it, _, _ = _cold(pm, 4, np.zeros((len(models), 2)), target=golden["tobi_rand_target"], per_instance=True)
    assert same_values(u0, golden["tobi_rand_u0"]) and same_values(ctr, golden["tobi_rand_controls"])
    assert np.array_equal(it, golden["tobi_rand_iters"])


@pytest.mark.parametrize("N", [10, 30, 60, 100])
def test_bike_cold_matches_dlib(golden, N):
    x0 = golden[f"bike{N}_cold_x0"]
    u0, ctr, it, st, _ = _cold(W.bicycle4_model(), N, x0)
    assert same_values(u0, golden[f"bike{N}_cold_u0"])
    assert same_values(ctr, golden[f"bike{N}_cold_controls"])
    assert np.array_equal(it, golden[f"bike{N}_cold_iters"])
    assert np.array_equal(st, (it >= 10000).astype(np.uint8))


def test_bike_closed_loop_with_moving_targets(golden):
    """Config 1: persistent controller, set_target(val, t) pre-fill, set_last_target each tick, 300 ticks."""
    bike, N = W.bicycle4_model(), 30
    with BatchedMPC(4, 1, N, 2) as eng:
        eng.set_model(bike)
        for t in range(N):
            eng.set_target([0.1 * np.sin(0.05 * t), 0, 0, 0], t)
        for k in range(300):
            eng.set_last_target([0.1 * np.sin(0.05 * (k + N)), 0, 0, 0])
            u0, it, _ = eng.solve(np.tile(golden["bike30_loop_x"][k], (2, 1)), warm=True)
            assert same_values(u0[0], golden["bike30_loop_u0"][k]) and same_values(u0[1], golden["bike30_loop_u0"][k]), k
            assert it[0] == golden["bike30_loop_iters"][k] == it[1], k


def test_sweep_slice_per_instance_models(golden):
    pm, x0 = W.sweep_models(per_combo=2, seed=44)
    u0, ctr, it, _, _ = _cold(pm, 10, x0, per_instance=True)
    assert same_values(u0, golden["sweep10_u0"]) and same_values(ctr, golden["sweep10_controls"])
    assert np.array_equal(it, golden["sweep10_iters"])


# ------------------------------------------------------------------------------------------ oracle differential
@pytest.mark.parametrize("shape", [(2, 2, 4), (2, 1, 30), (4, 1, 10), (4, 2, 10), (3, 2, 7), (4, 1, 30)])
def test_random_models_vs_oracle(oracle, shape):
    S, I, N = shape
    rng = np.random.default_rng(100 + S * 7 + I * 3 + N)
    B = 300
    pm = dict(A=np.eye(S) + 0.3 * rng.standard_normal((B, S, S)), B=rng.standard_normal((B, S, I)),
              C=0.05 * rng.standard_normal((B, S)), Q=rng.uniform(0, 10, (B, S)) * (rng.random((B, S)) > 0.25),
              R=rng.uniform(0.01, 2, (B, I)), lower=-rng.uniform(0.1, 1, (B, I)), upper=rng.uniform(0.1, 1, (B, I)))
    x0 = rng.standard_normal((B, S)); tgt = rng.standard_normal((B, S))
    with BatchedMPC(S, I, N, B) as eng:
        eng.set_model(pm, per_instance=True)
        eng.set_target(tgt)
        for b in (0, 17, B - 1):
            lam, qd = eng.derived(b)
            oc = oracle.controller({k: v[b] for k, v in pm.items()}, N)
            assert lam == oc.lam and same_values(qd, oc.q_diag())
        for eps, mi in ((0.01, 10000), (1e-6, 300), (0.5, 49), (0.01, 50), (0.01, 51), (0.01, 0)):
            eng.set_options(eps, mi)
            u0, it, st = eng.solve(x0, keep_controls=True)
            ctr = eng.get_controls(B)
            uo, co, ito, _ = oracle.batch_cold(pm, N, x0, target=tgt, eps=eps, max_iter=mi, per_instance=True, nthreads=4)
            assert same_values(u0, uo) and same_values(ctr, co), (shape, eps, mi)
            assert np.array_equal(it, ito), (shape, eps, mi)
            assert np.array_equal(st, (ito >= mi).astype(np.uint8))


@pytest.mark.parametrize("shape", [(2, 2, 4), (2, 1, 30), (4, 2, 10), (4, 1, 30)])
def test_random_models_symmetric_bounds_vs_oracle(oracle, shape):
    """Dense random models with lower == -upper: stage B runs its symmetric-bound kernel (pg_kernel.cuh, SYM), whose
    compares are reformulated (carried on-a-bound mask, one |x| < h test per clamp, integer-pipe compares).  Same bar
    as everywhere: value-exact controls, identical iteration counts — per-instance models, then one shared model."""
    S, I, N = shape
    rng = np.random.default_rng(500 + S * 7 + I * 3 + N)
    B = 400
    hi = rng.uniform(0.05, 1, (B, I))
    pm = dict(A=np.eye(S) + 0.25 * rng.standard_normal((B, S, S)), B=rng.standard_normal((B, S, I)),
              C=0.05 * rng.standard_normal((B, S)), Q=rng.uniform(0, 10, (B, S)) * (rng.random((B, S)) > 0.25),
              R=rng.uniform(0.01, 2, (B, I)), lower=-hi, upper=hi)
    x0 = rng.standard_normal((B, S)); tgt = rng.standard_normal((B, S))
    for eps, mi in ((0.01, 10000), (1e-5, 400), (0.01, 51)):
        uo, co, ito, _ = oracle.batch_cold(pm, N, x0, target=tgt, eps=eps, max_iter=mi, per_instance=True, nthreads=4)
        assert (ito > 50).sum() > B // 10, "too few instances reach the projected-gradient phase"
        for split in (1, 3):
            with BatchedMPC(S, I, N, B) as eng:
                eng.set_split(split)
                eng.set_model(pm, per_instance=True)
                eng.set_target(tgt)
                eng.set_options(eps, mi)
                u0, it, st = eng.solve(x0, keep_controls=True)
                assert same_values(u0, uo) and same_values(eng.get_controls(B), co), (shape, eps, mi, split)
                assert np.array_equal(it, ito) and np.array_equal(st, (ito >= mi).astype(np.uint8)), (shape, eps, mi, split)
    one = {k: v[7] for k, v in pm.items()}
    uo, co, ito, _ = oracle.batch_cold(one, N, x0, target=tgt, nthreads=4)
    with BatchedMPC(S, I, N, B) as eng:
        eng.set_model(one)
        eng.set_target(tgt)
        u0, it, _ = eng.solve(x0, keep_controls=True)
        assert same_values(u0, uo) and same_values(eng.get_controls(B), co) and np.array_equal(it, ito)


@pytest.mark.parametrize("split", [1, 0, 2, 3])
def test_cold_solves_leave_persistent_state_alone(oracle, split):
    """include/bmpc.h: a cold solve persists nothing.  warm -> cold (no KEEP_CONTROLS) -> warm on ONE handle must equal
    a controller object that only saw the two warm calls — in every bmpc_set_split setting (the two-stage solve used to
    park stage-A state in the persistent controls / v; it now parks in the hand-over ring's own slots)."""
    m, N, B = W.bicycle4_model(), 30, 96
    x0 = W.random_offsets_x0(B, 21)
    ocs = [oracle.controller(m, N) for _ in range(B)]
    with BatchedMPC(4, 1, N, B) as eng:
        eng.set_split(split)
        eng.set_model(m)
        u0, it, _ = eng.solve(x0, warm=True)
        uo = np.stack([ocs[b].step(x0[b]) for b in range(B)])
        assert same_values(u0, uo) and np.array_equal(it, [o.iters for o in ocs])
        assert (it > 50).mean() > 0.9
        ctr, = (eng.get_controls(B),)
        eng.solve(W.random_offsets_x0(B, 22))                       # cold, other inputs, nothing kept
        eng.solve(-x0[:40])                                         # cold and ragged
        assert np.array_equal(eng.get_controls(B).view(np.uint64), ctr.view(np.uint64))
        x1 = x0 @ m.A.T + u0 @ m.B.T
        u1, it1, _ = eng.solve(x1, warm=True)
        uo1 = np.stack([ocs[b].step(x1[b]) for b in range(B)])
        assert same_values(u1, uo1) and np.array_equal(it1, [o.iters for o in ocs])
        assert same_values(eng.get_controls(B), np.stack([o.controls() for o in ocs]))


def test_ragged_warm_solve_with_step_targets(oracle):
    """A BMPC_WARM solve of count < capacity on a handle with per-step targets: the controllers that were not solved are
    untouched dlib::mpc objects (mpc.h:236-237 runs only inside operator()), although the target ring's head is shared."""
    m, N, B, cnt = W.bicycle4_model(1.2, 12, 4, 0.02), 10, 40, 23
    rng = np.random.default_rng(9)
    ocs = [oracle.controller(m, N) for _ in range(B)]
    x = 0.3 * rng.standard_normal((B, 4))
    with BatchedMPC(4, 1, N, B) as eng:
        eng.set_model(m)
        for t in range(N):
            val = 0.1 * rng.standard_normal((B, 4))
            eng.set_target(val, t)
            for b in range(B):
                ocs[b].set_target(val[b], t)
        for k in range(5):
            n = cnt if k % 2 == 0 else B                              # ragged, full, ragged, ...
            u0, it, _ = eng.solve(x[:n], warm=True)
            uo = np.stack([ocs[b].step(x[b]) for b in range(n)])
            assert same_values(u0, uo) and np.array_equal(it, [ocs[b].iters for b in range(n)]), k
            for b in (0, cnt - 1, cnt, B - 1):
                for tt in (0, 3, N - 1):
                    assert same_values(eng.get_target(tt, instance=b), ocs[b].targets()[tt]), (k, b, tt)
            x[:n] = x[:n] @ m.A.T + u0 @ m.B.T


def test_nonfinite_inputs_are_flagged(oracle):
    """status bit BMPC_ST_NONFINITE: a NaN / inf in x0 or the targets (the reference logs isnan() of its outputs,
    src/trajectory_point_follower.cpp:100-103).  Neighbours are unaffected.  The flagged solves are NOT claimed to
    reproduce dlib's NaN propagation: dlib multiplies structural zeros out (0 * NaN = NaN, also inside diagm(Q) * x),
    the kernels skip those products — hence the flag; their controls still lie inside the box."""
    from trajectory_controller_b200.engine import ST_NONFINITE
    m, N, B = W.bicycle4_model(), 30, 256
    x0 = W.random_offsets_x0(B, 5)
    x0[3, 0] = np.nan; x0[77, 1] = np.inf; x0[200, 2] = -np.inf; x0[201, 3] = np.nan
    bad = np.zeros(B, dtype=bool); bad[[3, 77, 200, 201]] = True
    uo, _, ito, _ = oracle.batch_cold(m, N, x0, nthreads=4, want_controls=False)
    for split, special in ((1, True), (0, True), (3, True), (1, False), (0, False)):
        with BatchedMPC(4, 1, N, B) as eng:
            eng.set_split(split)
            eng.set_specialization(special)
            eng.set_model(m)
            u0, it, st = eng.solve(x0)
            assert np.all((st[bad] & ST_NONFINITE) != 0) and np.all((st[~bad] & ST_NONFINITE) == 0)
            assert same_values(u0[~bad], uo[~bad]) and np.array_equal(it[~bad], ito[~bad]), split
            assert np.all(np.abs(u0[bad]) <= W.ALPHA_MAX)          # whatever came out is inside the box and not a NaN


def test_model_validation():
    """dlib's requires clause (mpc.h:92-94) is enforced at the boundary instead of being undefined behaviour."""
    from trajectory_controller_b200.engine import BmpcError
    import copy
    good = W.bicycle4_model()
    with BatchedMPC(4, 1, 10, 4) as eng:
        for field, val in (("R", np.array([0.0])), ("Q", np.array([1.0, -1.0, 0, 0])), ("upper", np.array([-1.0])),
                           ("A", np.full((4, 4), np.nan)), ("lower", np.array([-np.inf]))):
            bad = copy.deepcopy(good); setattr(bad, field, val)
            with pytest.raises(BmpcError):
                eng.set_model(bad)
        eng.set_model(good)
        u0, it, st = eng.solve(np.zeros((4, 4)))
        assert np.all(u0 == 0) and np.all(it == 0)


@pytest.mark.parametrize("which", ["bike10", "kat"])
def test_persistent_controllers_vs_oracle(oracle, which):
    """Warm batch: per-instance states, per-step targets set at random times, shift semantics (mpc.h:229-237) and the
    stale-v corner (Q_diag == 0 at the 50th coordinate step, reachable with dlib's KAT model)."""
    rng = np.random.default_rng(5)
    if which == "bike10":
        m, N, B, ticks = W.bicycle4_model(1.3, 10, 3, 0.01), 10, 48, 40
    else:
        m, N, B, ticks = W.kat_model(), 30, 16, 12
    S = m.S
    ocs = [oracle.controller(m, N) for _ in range(B)]
    x = rng.standard_normal((B, S)) * (0.3 if which == "bike10" else 2.0)
    with BatchedMPC(S, m.I, N, B) as eng:
        eng.set_model(m)
        for k in range(ticks):
            if k % 3 == 0:
                t = int(rng.integers(0, N)); val = rng.standard_normal((B, S)) * 0.1
                eng.set_target(val, t)
                for b in range(B):
                    ocs[b].set_target(val[b], t)
            if k == 20:
                val = rng.standard_normal(S) * 0.05
                eng.set_target(val)                       # all steps, all instances
                for b in range(B):
                    ocs[b].set_target(val)
            u0, it, _ = eng.solve(x, warm=True)
            uo = np.stack([ocs[b].step(x[b]) for b in range(B)])
            assert same_values(u0, uo), k
            assert np.array_equal(it, [ocs[b].iters for b in range(B)]), k
            if k % 7 == 0:
                assert same_values(eng.get_controls(B), np.stack([ocs[b].controls() for b in range(B)]))
                for tt in (0, N // 2, N - 1):
                    assert same_values(eng.get_target(tt, instance=3), ocs[3].targets()[tt])
            x = x @ m.A.T + u0 @ m.B.T + 0.01 * rng.standard_normal((B, S))


# ------------------------------------------------------------------------------------------ full size, properties
def test_config2_full_batch_properties(oracle):
    """BASELINE config 2 at full size (65,536 cold instances, N=30, FP64): EVERY instance against the oracle (u0 value-
    exact, identical iteration counts; a few seconds on the host cores), then size-independent properties: box
    feasibility, determinism, batch-order independence.  The launch runs all three stages (stage B hands its last
    instances to the lane-pair kernel)."""
    import os
    bike, N, B = W.bicycle4_model(), 30, 65536
    x0 = W.random_offsets_x0(B, 42)
    with BatchedMPC(4, 1, N, B) as eng:
        eng.set_model(bike)
        u0, it, st = eng.solve(x0)
        uo, _, ito, _ = oracle.batch_cold(bike, N, x0, nthreads=os.cpu_count() or 8, want_controls=False)
        assert same_values(u0, uo) and np.array_equal(it, ito)
        assert np.all(np.abs(u0) <= W.ALPHA_MAX) and np.all(it <= 10000) and np.all(st == 0)
        u1, it1, _ = eng.solve(x0)                                      # determinism / no state leaks in cold mode
        assert np.array_equal(u0.view(np.uint64), u1.view(np.uint64)) and np.array_equal(it, it1)
        perm = np.random.default_rng(1).permutation(B)                  # which lane solves an instance is irrelevant
        u2, it2, _ = eng.solve(x0[perm])
        assert np.array_equal(u0[perm].view(np.uint64), u2.view(np.uint64)) and np.array_equal(it[perm], it2)
        # mirror symmetry of the problem: x0 -> -x0 gives u0 -> -u0 with the same iteration counts
        u3, it3, _ = eng.solve(-x0[:4096])
        assert same_values(u3, -u0[:4096]) and np.array_equal(it3, it[:4096])


def test_ragged_batches_and_capacity(oracle):
    """count < capacity, count = 1, count not a multiple of the warp size."""
    bike, N = W.bicycle4_model(), 10
    x0 = W.random_offsets_x0(1000, 7)
    uo, _, ito, _ = oracle.batch_cold(bike, N, x0)
    with BatchedMPC(4, 1, N, 1000) as eng:
        eng.set_model(bike)
        for cnt in (1, 31, 33, 257, 1000):
            u0, it, _ = eng.solve(x0[:cnt])
            assert same_values(u0, uo[:cnt]) and np.array_equal(it, ito[:cnt]), cnt


def test_device_results_block_layout():
    """u0 | iters | status of a handle are one device allocation (what a multi-GPU host gathers with one collective)."""
    import torch
    B = 5000
    x0 = W.random_offsets_x0(B, 3)
    with BatchedMPC(4, 1, 30, B) as eng:
        eng.set_model(W.bicycle4_model())
        u0, it, st = eng.solve(x0)
        view, off_it, off_st = eng.device_results_block()
        blk = torch.as_tensor(view, device="cuda").cpu().numpy()
        assert off_it % 256 == 0 and off_st % 256 == 0 and off_it >= 8 * B and off_st >= off_it + 4 * B and blk.size >= off_st + B
        assert np.array_equal(blk[: 8 * B].view(np.float64), u0[:, 0])
        assert np.array_equal(blk[off_it: off_it + 4 * B].view(np.uint32), it)
        assert np.array_equal(blk[off_st: off_st + B], st)
        du, di, ds = eng.device_results()
        assert np.array_equal(torch.as_tensor(di, device="cuda").cpu().numpy().view(np.uint32), it)


def test_error_paths_on_device():
    from trajectory_controller_b200.engine import BmpcError
    with BatchedMPC(4, 1, 10, 8) as eng:
        with pytest.raises(BmpcError):
            eng.solve(np.zeros((4, 4)))                 # no model yet
        eng.set_model(W.bicycle4_model())
        with pytest.raises(BmpcError):
            eng.solve(np.zeros((9, 4)))                 # count > capacity
        with pytest.raises(BmpcError):
            eng.set_target(np.zeros(4), 10)             # time >= horizon
        with pytest.raises(BmpcError):
            eng.set_options(0.0, 10)                    # eps must be > 0
        assert eng.get_epsilon() == 0.01 and eng.get_max_iterations() == 10000   # mpc.h:103-104 defaults


# ------------------------------------------------------------------------------------------ trajectory point
def test_extract_points_vs_oracle(oracle):
    rng = np.random.default_rng(3)
    B, P = 5000, 64
    px = np.cumsum(rng.uniform(0.01, 0.2, (B, P)), axis=1).astype(np.float32)
    py = (0.3 * np.sin(px) + 0.01 * rng.standard_normal((B, P))).astype(np.float32)
    dx, dy, vel = (rng.standard_normal((B, P)).astype(np.float32) for _ in range(3))
    dist = rng.uniform(-0.1, 8.0, B).astype(np.float32)
    npts = rng.integers(0, P + 1, B).astype(np.int32)
    with BatchedMPC(2, 2, 4, 8) as eng:
        for layout, arrs in (("bp", (px, py, dx, dy, vel)), ("pb", tuple(np.ascontiguousarray(a.T) for a in (px, py, dx, dy, vel)))):
            for n in (None, npts):
                gi, gf, go = eng.extract_points(*arrs, dist, npoints=n, layout=layout)
                oi, of, oo = oracle.extract(*arrs, dist, npoints=n, layout=layout)
                assert np.array_equal(gi, oi) and np.array_equal(gf, of)
                for a, b in zip(go, oo):
                    assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        # one shared polyline, many lookahead distances, incl. exact hits of the cumulative length (strict '>')
        sx = np.arange(P, dtype=np.float32) * np.float32(0.125); z = np.zeros(P, dtype=np.float32)
        d = np.concatenate([np.arange(0, 70, dtype=np.float32) * np.float32(0.125), rng.uniform(-1, 9, 500).astype(np.float32)])
        gi, gf, go = eng.extract_points(sx, z, z + 1, z, z + 2, d)
        oi, of, oo = oracle.extract(sx, z, z + 1, z, z + 2, d)
        assert np.array_equal(gi, oi) and np.array_equal(gf, of)
        for a, b in zip(go, oo):
            assert np.array_equal(a.view(np.uint32), b.view(np.uint32))
        # empty trajectory: default point (d,0),(1,0),0 with idx -1
        gi, gf, go = eng.extract_points(z[:0], z[:0], z[:0], z[:0], z[:0], np.float32([0.7]))
        assert gi[0] == -1 and gf[0] == 2 and go[0][0] == np.float32(0.7) and go[2][0] == 1.0


# ------------------------------------------------------------------------------------------ two-stage solve
def test_extract_horizon_vs_oracle(oracle):
    """bmpc_extract_horizon (one scan, N look-ahead distances) against the restatement, all three layouts, ragged and
    empty polylines, steps from 1 to 100 (shared-memory tile sizes), step 0."""
    rng = np.random.default_rng(21)
    B, P = 777, 37
    ang = np.cumsum(rng.normal(0, 0.15, (B, P)), axis=1)
    seg = rng.uniform(0.02, 0.3, (B, P))
    arr = [np.cumsum(seg * np.cos(ang), 1), np.cumsum(seg * np.sin(ang), 1), np.cos(ang), np.sin(ang), rng.uniform(0, 2, (B, P))]
    arr = [a.astype(np.float32) for a in arr]
    npts = rng.integers(0, P + 1, B).astype(np.int32); npts[:5] = [0, 1, 2, P, P]
    d0 = rng.uniform(-0.1, 1.5, B).astype(np.float32); step = rng.choice([0.0, 0.05, 0.1, 0.4], B).astype(np.float32)
    with BatchedMPC(4, 1, 30, B) as eng:
        for steps in (1, 30, 100):
            want = oracle.extract_horizon(*arr, d0, step, steps, npoints=npts, layout="bp")
            for layout, a in (("bp", arr), ("pb", [np.ascontiguousarray(x.T) for x in arr])):
                got = eng.extract_horizon(*a, d0, step, steps, npoints=npts, layout=layout)
                assert np.array_equal(got[0], want[0]) and np.array_equal(got[1], want[1]), (steps, layout)
                for g, w in zip(got[2], want[2]):
                    assert np.array_equal(g.view(np.uint32), w.view(np.uint32)), (steps, layout)
        one = [x[3] for x in arr]
        want = oracle.extract_horizon(*one, d0, step, 30)
        got = eng.extract_horizon(*one, d0, step, 30, layout="shared")
        assert np.array_equal(got[0], want[0])
        for g, w in zip(got[2], want[2]):
            assert np.array_equal(g.view(np.uint32), w.view(np.uint32))
        # step t of the horizon call == the single-point call at d_t
        t = 7
        dt = (d0 + np.float32(t) * step).astype(np.float32)
        i1, f1, o1 = eng.extract_points(*arr, dt, npoints=npts, layout="bp")
        got = eng.extract_horizon(*arr, d0, step, 30, npoints=npts, layout="bp")
        assert np.array_equal(got[0][:, t], i1) and np.array_equal(got[2][1][:, t].view(np.uint32), o1[1].view(np.uint32))
        with pytest.raises(Exception):
            eng.extract_horizon(*arr, d0, -np.abs(step) - 1, 30, layout="bp")


@pytest.mark.parametrize("shape", ["bike30", "bike30_pim", "bike60", "kat30", "tobi4", "dense427"])
def test_split_solve_equals_one_kernel(shape):
    """bmpc_set_split: stage A (set-up + coordinate phase) + stage B (projected-gradient kernel, pg_kernel.cuh) must
    reproduce the one-kernel solve bit for bit — u0, the whole control sequence, iteration counts, status — for cold
    and warm solves, shared and per-instance models, and when max_iterations cuts the projected-gradient phase."""
    rng = np.random.default_rng(11)
    if shape.startswith("bike30"):
        m, N, B, scale = W.bicycle4_model(), 30, 3000, None
    elif shape == "bike60":                                   # controls / linear term resident in shared memory in stage B
        m, N, B, scale = W.bicycle4_model(), 60, 700, None
    elif shape == "kat30":
        m, N, B, scale = W.kat_model(), 30, 500, 3.0
    elif shape == "tobi4":
        m, N, B, scale = W.tobi_model(), 4, 500, 0.5
    else:
        A = rng.standard_normal((3, 3)) * 0.5 + np.eye(3)
        m = W.Model(A, rng.standard_normal((3, 2)), rng.standard_normal(3) * 0.01, np.array([1.0, 2.0, 0.5]),
                    np.array([0.1, 0.2]), np.array([-0.4, -0.3]), np.array([0.5, 0.3]))
        N, B, scale = 7, 500, 1.0
    S, I = m.S, m.I
    x0 = W.random_offsets_x0(B, 3) if scale is None else rng.standard_normal((B, S)) * scale
    model, pim = m, False
    if shape == "bike30_pim":
        pim = True
        speeds = rng.uniform(0.5, 2.0, B)
        ms = [W.bicycle4_model(v=float(v)) for v in speeds]
        model = {k: np.stack([getattr(mm, k) for mm in ms]) for k in ("A", "B", "C", "Q", "R", "lower", "upper")}
    res = []
    for split in (1, 0, 2, 3):
        with BatchedMPC(S, I, N, B) as eng:
            eng.set_split(split)
            eng.set_model(model, per_instance=pim)
            out = []
            for max_iter in (10000, 120):
                eng.set_options(0.01 if shape != "kat30" else 1e-6, max_iter)
                u0, it, st = eng.solve(x0, keep_controls=True)
                out += [u0.copy(), it.copy(), st.copy(), eng.get_controls(B)]
            eng.set_options(0.01, 10000)
            eng.reset()
            x = x0.copy()
            for k in range(4):                               # persistent controllers (warm start, v carried over)
                u0, it, st = eng.solve(x, warm=True)
                out += [u0.copy(), it.copy(), st.copy(), eng.get_controls(B)]
                Am = m.A if not pim else model["A"]
                x = (np.einsum("bij,bj->bi", Am, x) if pim else x @ m.A.T) + (u0 @ m.B.T if not pim else np.einsum("bij,bj->bi", model["B"], u0))
            res.append(out)
    assert any((o > 50).any() for o in res[0][1::4]), "the projected-gradient stage never ran"
    for other in res[1:]:
        for a, b in zip(res[0], other):
            if a.dtype == np.float64:
                assert np.array_equal(a.view(np.uint64), b.view(np.uint64))
            else:
                assert np.array_equal(a, b)


# ------------------------------------------------------------------------------------------ tolerance modes
def test_tolerance_modes_fast_and_fp32(oracle):
    """BASELINE configs[2] in small: FP64 FAST (FMA allowed) and the two FP32 modes against the FP64 EXACT result (itself
    checked against the oracle here).  Stated tolerances (north star): |du0| <= 1e-6 rad in FP64 mode, <= 1e-3 rad in FP32.
      FP64 FAST : bar MET with seven orders of margin and identical iteration counts — this is the engine's fast mode.
      FP32      : bar NOT MET at the maximum (6.4e-3 over 1,048,576 instances, profiles/r2_cfg3_n*.json); met at the
                  99.9th percentile (5e-4).  The outliers are solves whose stopping test (max_df < eps, mpc.h:310) trips
                  dozens of iterations earlier or later than in FP64 while u0 is still moving: dlib's result is the
                  iterate at ITS trip point, not the optimum, so no FP32 trajectory can reproduce it to 1e-3 without
                  finishing in FP64.  The asserts below pin what FP32 does deliver (p99.9 <= 1e-3, max <= 2e-2, mean
                  <= 1e-4); README / DESIGN.md / bench.py report FP32 as failing the north-star bar."""
    from trajectory_controller_b200.engine import FAST, FP32
    bike, N, B = W.bicycle4_model(), 30, 20000
    x0 = W.random_offsets_x0(B, 43)
    res = {}
    for name, prec, mode in (("e64", FP64, EXACT), ("f64", FP64, FAST), ("e32", FP32, EXACT), ("f32", FP32, FAST)):
        with BatchedMPC(4, 1, N, B, prec, mode) as eng:
            eng.set_model(bike)
            u0, it, st = eng.solve(x0)
            res[name] = (u0[:, 0].copy(), it.astype(np.int64), st.copy())
    pick = np.arange(0, B, 40)
    uo, _, ito, _ = oracle.batch_cold(bike, N, x0[pick], nthreads=8, want_controls=False)
    assert same_values(res["e64"][0][pick], uo[:, 0]) and np.array_equal(res["e64"][1][pick], ito)
    ref_u, ref_it, _ = res["e64"]
    du = np.abs(res["f64"][0] - ref_u)
    assert du.max() <= 1e-6 and du.max() < 1e-11, du.max()
    assert (res["f64"][1] == ref_it).mean() > 0.999
    for name in ("e32", "f32"):
        u, it, st = res[name]
        du = np.abs(u - ref_u)
        assert np.percentile(du, 99.9) <= 1e-3, (name, np.percentile(du, 99.9))
        assert du.mean() <= 1e-4 and du.max() <= 2e-2, (name, du.mean(), du.max())
        assert np.all(np.abs(u) <= np.float32(W.ALPHA_MAX) * (1 + 1e-6)) and np.all(st == 0)
        assert abs(it.mean() - ref_it.mean()) < 5.0


def test_config3_one_million_instances_properties(oracle):
    """BASELINE configs[2] at this rank's full size (1,048,576 cold instances, seed 43, FP64 EXACT; stage B runs without
    time slicing at this size): 65,536 of them against the oracle, box feasibility, and the mirror symmetry x0 -> -x0."""
    bike, N, B = W.bicycle4_model(), 30, 1 << 20
    x0 = W.random_offsets_x0(B, 43)
    with BatchedMPC(4, 1, N, B) as eng:
        eng.set_model(bike)
        u0, it, st = eng.solve(x0)
        import os
        pick = np.random.default_rng(3).choice(B, 65536, replace=False)
        uo, _, ito, _ = oracle.batch_cold(bike, N, x0[pick], nthreads=os.cpu_count() or 8, want_controls=False)
        assert same_values(u0[pick], uo) and np.array_equal(it[pick], ito)
        assert np.all(np.abs(u0) <= W.ALPHA_MAX) and np.all(st == 0) and it.max() <= 10000
        u1, it1, _ = eng.solve(-x0)
        assert same_values(u1, -u0) and np.array_equal(it1, it)


@pytest.mark.parametrize("eps,max_iter", [(0.01, 10000), (0.01, 60), (0.001, 3000)])
def test_stage_thresholds_are_invisible(eps, max_iter):
    """Batch sizes around every scheduling threshold of the staged solve (stage C direct up to 9,472 instances; above it
    stage B with the hand-over of its last 9,472; time slices from 37,889; small max_iterations end instances inside every
    stage): bit-equal to stages A + B alone (bmpc_set_split 2), cold and over a warm sequence."""
    bike, cap = W.bicycle4_model(), 80000
    x0 = W.random_offsets_x0(cap, 7)
    with BatchedMPC(4, 1, 30, cap) as a, BatchedMPC(4, 1, 30, cap) as b:
        a.set_model(bike); b.set_model(bike); b.set_split(2)
        a.set_options(eps, max_iter); b.set_options(eps, max_iter)
        for n in (1, 33, 9471, 9473, 18945, 37889, 75777):
            ua, ia, sa = a.solve(x0[:n]); ub, ib, sb = b.solve(x0[:n])
            assert np.array_equal(ua.view(np.uint64), ub.view(np.uint64)) and np.array_equal(ia, ib) and np.array_equal(sa, sb), n
        n = 40000
        a.reset(); b.reset()
        x = x0[:n].copy()
        for tick in range(3):
            ua, ia, _ = a.solve(x, warm=True); ub, ib, _ = b.solve(x, warm=True)
            assert np.array_equal(ua.view(np.uint64), ub.view(np.uint64)) and np.array_equal(ia, ib), tick
            x = x * 0.8; x[:, 2] += 0.1 * ua[:, 0]
        assert np.array_equal(a.get_controls(n).view(np.uint64), b.get_controls(n).view(np.uint64))


def test_time_slicing_is_invisible():
    """Stage B parks and resumes instances every 128 iterations when a batch gives each lane 1-4 instances
    (bmpc_solve.cu).  65,536 and 100,000 instances with time slicing (split = 1), without (2) and on the one-kernel
    path (0): u0, the control sequences, iteration counts and status must agree bit for bit, cold and warm."""
    bike, N = W.bicycle4_model(), 30
    for B in (65536, 100000):
        x0 = W.random_offsets_x0(B, 7)
        res = []
        for split in (1, 2, 0, 3):
            with BatchedMPC(4, 1, N, B) as eng:
                eng.set_split(split)
                eng.set_model(bike)
                u0, it, st = eng.solve(x0, keep_controls=True)
                out = [u0.copy(), it.copy(), st.copy(), eng.get_controls(B)]
                eng.reset()
                u0, it, st = eng.solve(x0, warm=True)
                u1, it1, st1 = eng.solve(x0 @ bike.A.T + u0 @ bike.B.T, warm=True)
                res.append(out + [u0.copy(), it.copy(), u1.copy(), it1.copy(), eng.get_controls(B)])
        assert (res[0][1] > 178).mean() > 0.5          # most instances were parked at least once
        for other in res[1:]:
            for a, b in zip(res[0], other):
                assert np.array_equal(a.view(np.uint64) if a.dtype == np.float64 else a, b.view(np.uint64) if b.dtype == np.float64 else b)
