"""Synthetic workloads for the batched MPC engine: models, seeds and input generators.

These are the concrete benchmark shapes of BASELINE.json ``configs`` as fixed by SURVEY.md §8(d).
Nothing here is reference behaviour except the constants that are cited; the (4,1,N) vehicle model is
survey-defined (the reference hard-codes a 2-state / 2-input model with horizon 4,
/root/reference/src/trajectory_point_follower.cpp:303-366).

Pure numpy; imported by tests/, bench.py and the examples.  No oracle and no CUDA in here.
"""
from __future__ import annotations

import math
from dataclasses import dataclass

import numpy as np

# /root/reference/src/trajectory_point_follower.cpp:16 (alpha_max = 22 deg), :96 (stepSize), header :47 (l)
ALPHA_MAX = 22 * math.pi / 180
STEP_T = 0.1
WHEELBASE = 0.21


@dataclass
class Model:
    """Row-major A (S,S), B (S,I), C (S,), Q (S,), R (I,), lower (I,), upper (I,) — dlib::mpc's 7 ctor args
    (/root/reference/dlib_files/dlib/control/mpc.h:51-59)."""
    A: np.ndarray
    B: np.ndarray
    C: np.ndarray
    Q: np.ndarray
    R: np.ndarray
    lower: np.ndarray
    upper: np.ndarray

    @property
    def S(self) -> int:
        return self.A.shape[-1]

    @property
    def I(self) -> int:
        return self.B.shape[-1]


def bicycle4_model(v: float = 1.0, w_y: float = 20.0, w_psi: float = 7.0, r: float = 0.05,
                   T: float = STEP_T, l: float = WHEELBASE, omega: float = 5.0, zeta: float = 1.0) -> Model:
    """The synthetic 4-state / 1-input lateral model of SURVEY.md §8(d): x=(e_y, e_psi, delta, delta_dot)."""
    A = np.array([[1, T * v, 0, 0],
                  [0, 1, T * v / l, 0],
                  [0, 0, 1, T],
                  [0, 0, -T * omega * omega, 1 - 2 * zeta * omega * T]], dtype=np.float64)
    B = np.array([[0], [0], [0], [T * omega * omega]], dtype=np.float64)
    return Model(A, B, np.zeros(4), np.array([w_y, w_psi, 0.0, 0.0]), np.array([r]),
                 np.array([-ALPHA_MAX]), np.array([ALPHA_MAX]))


def tobi_model(v: float = 1.0, w_y: float = 20.0, w_phi: float = 7.0, w_sf: float = 0.0005, w_sr: float = 10.0,
               T: float = STEP_T, l: float = WHEELBASE) -> Model:
    """The reference's own model, mpcControllerTobi (/root/reference/src/trajectory_point_follower.cpp:326-363)."""
    A = np.array([[1, T * v], [0, 1]], dtype=np.float64)
    B = np.array([[0, T * v], [T * v / l, -T * v / l]], dtype=np.float64)
    return Model(A, B, np.zeros(2), np.array([w_y, w_phi]), np.array([w_sf, w_sr]),
                 np.array([-ALPHA_MAX, -ALPHA_MAX]), np.array([ALPHA_MAX, ALPHA_MAX]))


def kat_model() -> Model:
    """dlib's own known-answer test model (/root/reference/dlib_files/dlib/test/mpc.cpp:269-293)."""
    A = np.array([[1, 1], [0, 1]], dtype=np.float64)
    B = np.array([[0], [1]], dtype=np.float64)
    return Model(A, B, np.array([0.02, 0.1]), np.array([2.0, 0.0]), np.array([1.0]),
                 np.array([-0.2]), np.array([0.2]))


class MT19937_64:
    """std::mt19937_64 + libstdc++'s std::uniform_real_distribution<double> (one 64-bit draw per value:
    generate_canonical<double,53> = double(x) / 2^64, clamped below 1), so that seeds 42-45 of SURVEY.md
    §8(d) mean the same numbers here and in a C++ harness."""
    NN, MM = 312, 156
    _UM, _LM = np.uint64(0xFFFFFFFF80000000), np.uint64(0x7FFFFFFF)
    _MAT = np.uint64(0xB5026F5AA96619E9)

    def __init__(self, seed: int):
        mt = [0] * self.NN
        mt[0] = seed & 0xFFFFFFFFFFFFFFFF
        for i in range(1, self.NN):
            mt[i] = (6364136223846793005 * (mt[i - 1] ^ (mt[i - 1] >> 62)) + i) & 0xFFFFFFFFFFFFFFFF
        self.mt = np.array(mt, dtype=np.uint64)
        self.buf = np.empty(0, dtype=np.uint64)

    def _twist(self, cur, nxt, far):
        x = (cur & self._UM) | (nxt & self._LM)
        return far ^ (x >> np.uint64(1)) ^ np.where(x & np.uint64(1), self._MAT, np.uint64(0))

    def _block(self) -> np.ndarray:
        mt, NN, MM = self.mt, self.NN, self.MM
        mt[:NN - MM] = self._twist(mt[:NN - MM], mt[1:NN - MM + 1], mt[MM:NN])
        mt[NN - MM:NN - 1] = self._twist(mt[NN - MM:NN - 1], mt[NN - MM + 1:NN], mt[:MM - 1])
        mt[NN - 1:] = self._twist(mt[NN - 1:], mt[:1], mt[MM - 1:MM])
        x = mt.copy()
        x ^= (x >> np.uint64(29)) & np.uint64(0x5555555555555555)
        x ^= (x << np.uint64(17)) & np.uint64(0x71D67FFFEDA60000)
        x ^= (x << np.uint64(37)) & np.uint64(0xFFF7EEE000000000)
        x ^= x >> np.uint64(43)
        return x

    def raw(self, n: int) -> np.ndarray:
        parts = [self.buf]
        have = len(self.buf)
        while have < n:
            b = self._block()
            parts.append(b)
            have += len(b)
        allv = np.concatenate(parts)
        self.buf = allv[n:]
        return allv[:n]

    def uniform(self, n: int, a: float = -1.0, b: float = 1.0) -> np.ndarray:
        r = self.raw(n).astype(np.float64) / 18446744073709551616.0
        r = np.where(r >= 1.0, np.nextafter(1.0, 0.0), r)
        return r * (b - a) + a


def random_offsets_x0(count: int, seed: int, S: int = 4) -> np.ndarray:
    """e_y = 0.3 U, e_psi = 0.4 U drawn in instance order (e_y then e_psi),
    remaining states 0.  Returns AoS [count][S] float64."""
    u = MT19937_64(seed).uniform(2 * count).reshape(count, 2)
    x0 = np.zeros((count, S), dtype=np.float64)
    x0[:, 0] = 0.3 * u[:, 0]
    x0[:, 1] = 0.4 * u[:, 1]
    return x0


def sweep_models(N_unused: int = 0, per_combo: int = 1024, seed: int = 44):
    """4x4x4 weight grid x `per_combo` random x0 each -> per-instance models.
    Returns (dict of per-instance arrays A[B,4,4] B[B,4,1] C Q R lower upper, x0[B,4])."""
    combos = [(wy, wp, r) for wy in (5.0, 10.0, 20.0, 40.0) for wp in (1.75, 3.5, 7.0, 14.0)
              for r in (5e-4, 5e-3, 5e-2, 5e-1)]
    count = len(combos) * per_combo
    arrs = {k: [] for k in ("A", "B", "C", "Q", "R", "lower", "upper")}
    for wy, wp, r in combos:
        m = bicycle4_model(1.0, wy, wp, r)
        for k in arrs:
            arrs[k].append(np.broadcast_to(getattr(m, k), (per_combo,) + getattr(m, k).shape))
    out = {k: np.ascontiguousarray(np.concatenate(v, axis=0)) for k, v in arrs.items()}
    return out, random_offsets_x0(count, seed)


def flops_per_iteration(S: int, I: int, N: int) -> int:
    """Algorithmic flops of one solver iteration, SURVEY.md §8(d) (dense A,B; diagonal Q,R; PG update)."""
    return (S * (2 * I - 1) + (N - 1) * S * (2 * S + 2 * I - 1) + N * S + (N - 1) * 2 * S * S
            + N * I * (2 * S + 2) + 5 * N * I)


def flops_setup(S: int, I: int, N: int) -> int:
    """Algorithmic flops of the per-solve linear-term setup (mpc.h:258-266), SURVEY.md §8(d)."""
    return N * 2 * S * S + N * 2 * S + (N - 1) * 2 * S * S + N * I * (2 * S - 1)


def flops_ctor(S: int, I: int, N: int) -> int:
    """Algorithmic flops of the constructor recursion (mpc.h:114-123), SURVEY.md §8(d)."""
    return N * (I * S * (2 * S - 1) + I * (2 * S - 1) + I + 2 * S * S * (2 * S - 1) + S)


def sine_world_path(length_m: float = 70.0, ds: float = 0.1, amp: float = 0.3, period: float = 4.0):
    """world path y(x) = amp*sin(2*pi*x/period) as a polyline with points every `ds`
    metres of ARC LENGTH (positions and unit tangents, float64).  Returns (x, y, tx, ty)."""
    k = 2 * math.pi / period
    fine = np.linspace(0.0, length_m, int(length_m / ds) * 200 + 1)
    dy = amp * k * np.cos(k * fine)
    seg = np.sqrt(1.0 + dy * dy)
    s = np.concatenate([[0.0], np.cumsum(0.5 * (seg[1:] + seg[:-1]) * np.diff(fine))])
    marks = np.arange(0.0, s[-1], ds)
    x = np.interp(marks, s, fine)
    y = amp * np.sin(k * x)
    t = amp * k * np.cos(k * x)
    n = np.sqrt(1.0 + t * t)
    return x, y, 1.0 / n, t / n


def closed_loop_poses(count: int, seed: int = 45, path=None):
    """Initial vehicle poses [count][4] = X, Y, cos(psi), sin(psi): lateral offset 0.3 U and heading offset 0.4 U
    (drawn per vehicle in that order, like random_offsets_x0) around the start of the path."""
    path = sine_world_path() if path is None else path
    u = MT19937_64(seed).uniform(2 * count).reshape(count, 2)
    psi0 = math.atan2(path[3][0], path[2][0])
    psi = psi0 + 0.4 * u[:, 1]
    # offset along the path normal at the start
    nx, ny = -path[3][0], path[2][0]
    pose = np.empty((count, 4))
    pose[:, 0] = path[0][0] + 0.3 * u[:, 0] * nx
    pose[:, 1] = path[1][0] + 0.3 * u[:, 0] * ny
    pose[:, 2] = np.cos(psi)
    pose[:, 3] = np.sin(psi)
    return pose
