"""Pins the CPU checkers (oracle/) before anything is compared against them.   [-m "not gpu"]

* oracle/mpc_oracle.c  vs the golden fixture generated from the real dlib::mpc (tests/golden/make_golden.py)
* oracle/mpc_oracle.c  vs oracle/_ref/libdlibref.so (the real header) on fresh random cases, where it exists
* both                 vs dlib's own known-answer test restated in numpy (dlib/test/mpc.cpp:24-129,178-318)
* oracle/extract_oracle.c vs a line-by-line numpy-float32 restatement of
  /root/reference/src/trajectory_point_follower.cpp:392-443 and its edge cases
"""
import numpy as np
import pytest

from _oracle import DlibRef, same_bits
from trajectory_controller_b200 import workloads as W


# ------------------------------------------------------------------------------------------------ solver
def test_rng_matches_libstdcxx(golden):
    assert same_bits(golden["rng42_head"], W.MT19937_64(42).uniform(700))


@pytest.mark.parametrize("tag,eps", [("kat_tight", 1e-8), ("kat_default", 0.01)])
def test_oracle_kat_closed_loop_bit_exact(oracle, golden, tag, eps):
    kat = W.kat_model()
    c = oracle.controller(kat, 30)
    c.set_options(eps, 10000)
    assert c.lam == golden[tag + "_lambda"][0]
    assert same_bits(c.q_diag(), golden[tag + "_qdiag"])
    x = np.array([5.0, 0.0])
    for k in range(30):
        assert same_bits(x, golden[tag + "_x"][k])
        u = c.step(x)
        assert same_bits(u, golden[tag + "_u0"][k]), k
        assert c.iters == golden[tag + "_iters"][k]
        x = np.array([(kat.A[0, 0] * x[0] + kat.A[0, 1] * x[1] + kat.B[0, 0] * u[0]) + kat.C[0],
                      (kat.A[1, 0] * x[0] + kat.A[1, 1] * x[1] + kat.B[1, 0] * u[0]) + kat.C[1]])
    if tag == "kat_default":
        assert golden[tag + "_u0"][29, 0] == -0.10000002376964306


def _dense_smo(Qm, b, alpha, lo, hi, eps=1e-8, max_iter=100000):
    """dlib/test/mpc.cpp:24-129 (solve_qp_box_using_smo) restated in numpy."""
    alpha = alpha.copy()
    df = Qm @ alpha + b
    qq = np.where(np.diag(Qm) != 0, 1.0 / np.where(np.diag(Qm) != 0, np.diag(Qm), 1.0), 0.0)
    for _ in range(max_iter):
        free = ~(((alpha <= lo) & (df > 0)) | ((alpha >= hi) & (df < 0)))
        a = np.where(free, np.abs(df), 0.0)
        r = int(np.argmax(a))
        max_df = a[r]
        old = alpha[r]
        alpha[r] = min(max(-(df[r] - Qm[r, r] * alpha[r]) * qq[r], lo[r]), hi[r])
        df -= Qm[r] * (old - alpha[r])
        if max_df < eps:
            break
    return alpha


def test_kat_against_independent_dense_qp(oracle, golden):
    """dlib/test/mpc.cpp:257-318: 30 ticks of mpc<2,1,30> agree with a condensed box-QP solved by SMO to 1e-7."""
    kat = W.kat_model()
    N, S = 30, 2
    Apow = [np.eye(2)]
    for _ in range(1, N):
        Apow.append(kat.A @ Apow[-1])
    K = np.zeros((N * S, N))
    for r in range(N):
        for c in range(r + 1):
            K[r * S:(r + 1) * S, c:c + 1] = Apow[r - c] @ kat.B
    QQ = np.kron(np.eye(N), np.diag(kat.Q))
    m1 = K.T @ QQ @ K + np.eye(N) * kat.R[0]
    lo, hi = np.full(N, kat.lower[0]), np.full(N, kat.upper[0])
    controls = np.zeros(N)
    for k in range(8):                    # the first 8 of dlib's 30 ticks (numpy SMO is slow; needs the warm-start chain)
        x = golden["kat_tight_x"][k]
        M = np.zeros(N * S)
        M[:S] = kat.A @ x + kat.C
        for i in range(1, N):
            M[i * S:(i + 1) * S] = kat.A @ M[(i - 1) * S:i * S] + kat.C
        m2 = K.T @ QQ @ M
        controls[:-1] = controls[1:]
        controls = _dense_smo(m1, m2, controls, lo, hi)
        assert abs(golden["kat_tight_u0"][k, 0] - controls[0]) < 1e-7, k


def test_oracle_tobi_spot_and_random(oracle, golden):
    u, _, _, _ = oracle.batch_cold(W.tobi_model(1.0), 4, np.zeros((1, 2)), target=np.array([[0.05, 0.1]]))
    assert tuple(u[0]) == (0.27357241694602841, 0.023830246920932288)
    assert same_bits(u[0], golden["tobi_spot_u0"])
    models = [W.tobi_model(float(v)) for v in golden["tobi_rand_v"]]
    pm = {k: np.stack([getattr(m, k) for m in models]) for k in ("A", "B", "C", "Q", "R", "lower", "upper")}
    u, c, it, _ = oracle.batch_cold(pm, 4, np.zeros((len(models), 2)), target=golden["tobi_rand_target"],
                                    per_instance=True)
    assert same_bits(u, golden["tobi_rand_u0"]) and same_bits(c, golden["tobi_rand_controls"])
    assert np.array_equal(it, golden["tobi_rand_iters"])


@pytest.mark.parametrize("N", [10, 30, 60, 100])
def test_oracle_bike_cold(oracle, golden, N):
    x0 = golden[f"bike{N}_cold_x0"]
    u, c, it, _ = oracle.batch_cold(W.bicycle4_model(), N, x0, nthreads=4)
    assert same_bits(u, golden[f"bike{N}_cold_u0"])
    assert same_bits(c, golden[f"bike{N}_cold_controls"])
    assert np.array_equal(it, golden[f"bike{N}_cold_iters"])


def test_oracle_bike_closed_loop(oracle, golden):
    bike, N = W.bicycle4_model(), 30
    c = oracle.controller(bike, N)
    for t in range(N):
        c.set_target([0.1 * np.sin(0.05 * t), 0, 0, 0], t)
    for k in range(300):
        c.set_target([0.1 * np.sin(0.05 * (k + N)), 0, 0, 0], N - 1)
        u = c.step(golden["bike30_loop_x"][k])
        assert same_bits(u, golden["bike30_loop_u0"][k]), k
        assert c.iters == golden["bike30_loop_iters"][k]


def test_oracle_sweep_slice(oracle, golden):
    pm, x0 = W.sweep_models(per_combo=2, seed=44)
    u, c, it, _ = oracle.batch_cold(pm, 10, x0, per_instance=True)
    assert same_bits(u, golden["sweep10_u0"]) and same_bits(c, golden["sweep10_controls"])
    assert np.array_equal(it, golden["sweep10_iters"])


def test_oracle_iteration_semantics(oracle):
    """mpc.h:270-271,310-311: iter==0 when already converged; loop ends at max_iterations otherwise."""
    bike = W.bicycle4_model()
    u, c, it, _ = oracle.batch_cold(bike, 30, np.zeros((1, 4)))
    assert it[0] == 0 and np.all(c == 0)
    u, c, it, _ = oracle.batch_cold(bike, 30, np.array([[0.3, 0.4, 0, 0]]), max_iter=7)
    assert it[0] == 7
    u, c, it, _ = oracle.batch_cold(bike, 30, np.array([[0.3, 0.4, 0, 0]]), max_iter=0)
    assert it[0] == 0 and np.all(c == 0)


@pytest.mark.ref
def test_oracle_vs_real_dlib_random(oracle, dlibref):
    """Fresh random differential against the real header: models, targets, eps, max_iter, warm starts."""
    rng = np.random.default_rng(7)
    for (S, I, N) in [(2, 2, 4), (2, 1, 30), (4, 1, 10), (4, 2, 10), (3, 2, 7), (4, 1, 30)]:
        B = 200
        pm = dict(A=np.eye(S) + 0.3 * rng.standard_normal((B, S, S)), B=rng.standard_normal((B, S, I)),
                  C=0.05 * rng.standard_normal((B, S)), Q=rng.uniform(0, 10, (B, S)) * (rng.random((B, S)) > 0.25),
                  R=rng.uniform(0.01, 2, (B, I)), lower=-rng.uniform(0.1, 1, (B, I)),
                  upper=rng.uniform(0.1, 1, (B, I)))
        x0 = rng.standard_normal((B, S)); tgt = rng.standard_normal((B, S))
        for eps, mi in ((0.01, 10000), (1e-6, 300), (0.5, 49), (0.01, 50), (0.01, 51)):
            u, c, _ = dlibref.batch_cold(pm, N, x0, target=tgt, eps=eps, max_iter=mi, per_instance=True, nthreads=4)
            uo, co, it, _ = oracle.batch_cold(pm, N, x0, target=tgt, eps=eps, max_iter=mi, per_instance=True,
                                              nthreads=4)
            assert same_bits(u, uo) and same_bits(c, co), (S, I, N, eps, mi)
            assert it.max() <= mi


@pytest.mark.ref
def test_oracle_vs_real_dlib_persistent(oracle, dlibref):
    """Persistent controller with per-step targets: shift semantics of mpc.h:229-239 incl. v[] carry-over."""
    rng = np.random.default_rng(11)
    m = W.bicycle4_model(1.3, 10, 3, 0.01)
    rc, oc = dlibref.controller(m, 10), oracle.controller(m, 10)
    x = np.array([0.25, -0.3, 0.05, 0.0])
    for k in range(80):
        if k % 3 == 0:
            t, val = int(rng.integers(0, 10)), rng.standard_normal(4) * 0.1
            rc.set_target(val, t); oc.set_target(val, t)
        u, uo = rc.step(x), oc.step(x)
        assert same_bits(u, uo) and same_bits(rc.controls(), oc.controls()) and same_bits(rc.targets(), oc.targets())
        x = m.A @ x + m.B @ u + 0.01 * rng.standard_normal(4)


# -------------------------------------------------------------------------------------------- extraction
def _extract_numpy(px, py, dx, dy, vel, d):
    """src/trajectory_point_follower.cpp:392-443 in numpy float32 scalars (every op rounds to float)."""
    f = np.float32
    P = len(px)
    if P == 0:
        return -1, 2, (f(d), f(0), f(1), f(0), f(0))
    cur = f(0)
    for i in range(1, P):
        ex, ey = f(px[i - 1] - px[i]), f(py[i - 1] - py[i])
        ln = np.sqrt(f(f(ex * ex) + f(ey * ey)))
        cur = f(cur + ln)
        if cur > d:
            delta = f(cur - d)
            ax, ay = f(f(ex / ln) * delta), f(f(ey / ln) * delta)
            return i, 0, (f(px[i] + ax), f(py[i] + ay), dx[i], dy[i], vel[i])
    return P - 1, 1, (px[-1], py[-1], dx[-1], dy[-1], vel[-1])


def test_extract_oracle_matches_numpy_restatement(oracle):
    rng = np.random.default_rng(3)
    for P in (0, 1, 2, 3, 17, 64):
        for _ in range(40):
            px = np.cumsum(rng.uniform(0.01, 0.2, P)).astype(np.float32)
            py = (0.3 * np.sin(px) + 0.01 * rng.standard_normal(P)).astype(np.float32)
            dx, dy, vel = (rng.standard_normal(P).astype(np.float32) for _ in range(3))
            for d in np.float32([0.0, -1.0, 0.05, 0.6, 1.7, 100.0]):
                idx, flag, outs = oracle.extract(px, py, dx, dy, vel, np.array([d], dtype=np.float32))
                ei, ef, eo = _extract_numpy(px, py, dx, dy, vel, d)
                assert (idx[0], flag[0]) == (ei, ef), (P, d)
                got = np.array([o[0] for o in outs], dtype=np.float32)
                assert np.array_equal(got.view(np.uint32), np.array(eo, dtype=np.float32).view(np.uint32)), (P, d)


def test_extract_oracle_edges(oracle):
    z = np.zeros(4, dtype=np.float32)
    px = np.array([0, 1, 2, 3], dtype=np.float32)
    one = np.ones(4, dtype=np.float32)
    # strict '>' at :428 — cumulative distance exactly equal to d does not select the segment
    idx, flag, outs = oracle.extract(px, z, one, z, one, np.array([1.0, 3.0, 2.999], dtype=np.float32))
    assert list(idx) == [2, 3, 3] and list(flag) == [0, 1, 0]
    assert outs[0][0] == 1.0 and outs[0][1] == 3.0
    # duplicate points: zero-length segment never satisfies cur > d for d >= 0, negative d hits it with 0/0 = NaN
    pxd = np.array([0, 0, 1, 2], dtype=np.float32)
    idx, flag, outs = oracle.extract(pxd, z, one, z, one, np.array([0.0, -0.5], dtype=np.float32))
    assert list(idx) == [2, 1] and np.isnan(outs[0][1])
    # per-instance point counts (ragged) and the three layouts agree
    rng = np.random.default_rng(5)
    B, P = 9, 6
    arr = [rng.standard_normal((B, P)).astype(np.float32) for _ in range(5)]
    arr[0] = np.cumsum(np.abs(arr[0]), axis=1).astype(np.float32)
    npts = np.array([0, 1, 2, 3, 4, 5, 6, 6, 6], dtype=np.int32)
    d = rng.uniform(0, 3, B).astype(np.float32)
    a = oracle.extract(*arr, d, npoints=npts, layout="bp")
    b = oracle.extract(*[np.ascontiguousarray(x.T) for x in arr], d, npoints=npts, layout="pb")
    assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1])
    for o1, o2 in zip(a[2], b[2]):
        assert np.array_equal(o1.view(np.uint32), o2.view(np.uint32))
    assert a[0][0] == -1 and a[1][0] == 2 and a[2][0][0] == d[0]


def test_extract_horizon_is_n_independent_lookups(oracle):
    """oracle/extract_oracle.c: the one-scan horizon extraction equals getTrajectoryPoint asked once per step with
    d_t = d0 + t*step (float ops rounded separately) — incl. several steps in one segment, the last-point fallback,
    empty and one-point polylines, step 0 and a NaN distance."""
    rng = np.random.default_rng(11)
    for trial in range(60):
        P = int(rng.integers(0, 40))
        ang = np.cumsum(rng.normal(0, 0.2, max(P, 1)))
        seg = rng.uniform(0.01, 0.5, max(P, 1))
        px = np.cumsum(seg * np.cos(ang)).astype(np.float32)[:P]; py = np.cumsum(seg * np.sin(ang)).astype(np.float32)[:P]
        dx = np.cos(ang).astype(np.float32)[:P]; dy = np.sin(ang).astype(np.float32)[:P]
        vel = rng.uniform(0, 2, max(P, 1)).astype(np.float32)[:P]
        d0 = np.float32(rng.uniform(-0.2, 2.0)); step = np.float32(rng.choice([0.0, 0.03, 0.1, 0.37, 1.5]))
        if trial == 7:
            d0 = np.float32(np.nan)
        steps = int(rng.integers(1, 40))
        if P == 0:
            z = np.zeros(1, dtype=np.float32)
            idx, flag, outs = oracle.extract_horizon(z, z, z, z, z, [d0], [step], steps, npoints=[0])
        else:
            idx, flag, outs = oracle.extract_horizon(px, py, dx, dy, vel, [d0], [step], steps)
        for t in range(steps):
            d = np.float32(d0 + np.float32(np.float32(t) * step))
            if P == 0:
                i1, f1, o1 = oracle.extract(z, z, z, z, z, np.array([d], dtype=np.float32), npoints=[0])
            else:
                i1, f1, o1 = oracle.extract(px, py, dx, dy, vel, np.array([d], dtype=np.float32))
            assert idx[0, t] == i1[0] and flag[0, t] == f1[0], (trial, t)
            for a, b in zip(outs, o1):
                assert a[0, t].view(np.uint32) == b[0].view(np.uint32) or (np.isnan(a[0, t]) and np.isnan(b[0])), (trial, t)
