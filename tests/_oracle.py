"""ctypes bindings for the CPU checkers under oracle/ (TEST INFRASTRUCTURE).

* ``Oracle``   -> oracle/liboracle.so   (plain-C restatement; always buildable; travels to the GPU box)
* ``DlibRef``  -> oracle/_ref/libdlibref.so (the real dlib::mpc; built only where /root/reference exists)

Only tests/, __graft_entry__.smoke() and bench.py's CPU legs import this module.
"""
from __future__ import annotations

import ctypes as C
import os
import subprocess

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
ORACLE_DIR = os.path.join(ROOT, "oracle")
_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)


def _d(a):
    return None if a is None else a.ctypes.data_as(_dp)


def _f(a):
    return None if a is None else a.ctypes.data_as(_fp)


def build_oracle(ref: bool = True) -> None:
    """(Re)build the checkers with oracle/Makefile.  `ref` targets are skipped when /root/reference is absent."""
    subprocess.run(["make", "-s", "-C", ORACLE_DIR, "all" if ref else os.path.join(ORACLE_DIR, "liboracle.so")],
                   check=True)


def _c64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


class _ModelArgs:
    """Flattens a workloads.Model (or per-instance dict of arrays) into the 7 pointer arguments."""

    def __init__(self, model, per_instance: bool = False):
        g = (lambda k: model[k]) if isinstance(model, dict) else (lambda k: getattr(model, k))
        self.arrs = [_c64(g(k)) for k in ("A", "B", "C", "Q", "R", "lower", "upper")]
        self.per_instance = per_instance
        if per_instance:
            self.S, self.I = self.arrs[1].shape[-2], self.arrs[1].shape[-1]
        else:
            self.S, self.I = self.arrs[1].shape[0], self.arrs[1].shape[1]

    @property
    def ptrs(self):
        return [_d(a) for a in self.arrs]


class Oracle:
    """The plain-C restatement (oracle/mpc_oracle.c, oracle/extract_oracle.c)."""

    def __init__(self):
        path = os.path.join(ORACLE_DIR, "liboracle.so")
        if not os.path.exists(path):
            build_oracle(ref=False)
        self.lib = L = C.CDLL(path)
        L.orc_batch_cold.restype = C.c_double
        L.orc_batch_cold.argtypes = [C.c_int] * 4 + [_dp] * 7 + [C.c_long, _dp, _dp, C.c_double, C.c_ulong,
                                                                _dp, _dp, C.POINTER(C.c_uint), C.c_int]
        L.orc_mpc_init.restype = C.c_int
        L.orc_mpc_init.argtypes = [C.c_void_p, C.c_int, C.c_int, C.c_int] + [_dp] * 7
        L.orc_mpc_set_target.argtypes = [C.c_void_p, _dp, C.c_long]
        L.orc_mpc_step.argtypes = [C.c_void_p, _dp, _dp]
        L.orc_extract_batch.argtypes = [C.c_long, C.c_int, C.POINTER(C.c_int32)] + [_fp] * 5 + \
            [C.c_long, C.c_long, _fp, C.POINTER(C.c_int32), C.POINTER(C.c_uint8)] + [_fp] * 5

    # ---- batched cold solves -------------------------------------------------------------------------
    def batch_cold(self, model, N, x0, target=None, eps=0.01, max_iter=10000, per_instance=False,
                   nthreads=1, want_controls=True):
        ma = _ModelArgs(model, per_instance)
        x0 = _c64(x0)
        B = x0.shape[0]
        tgt = None if target is None else _c64(target)
        u0 = np.zeros((B, ma.I)); ctr = np.zeros((B, N, ma.I)) if want_controls else None
        iters = np.zeros(B, dtype=np.uint32)
        secs = self.lib.orc_batch_cold(ma.S, ma.I, N, int(per_instance), *ma.ptrs, B, _d(x0), _d(tgt), eps,
                                       max_iter, _d(u0), _d(ctr), iters.ctypes.data_as(C.POINTER(C.c_uint)),
                                       nthreads)
        if secs < 0:
            raise ValueError("oracle: shape out of bounds")
        return u0, ctr, iters, secs

    # ---- one persistent controller ----------------------------------------------------------------------
    def controller(self, model, N):
        return OracleController(self, model, N)

    # ---- closed loop (oracle/closed_loop_oracle.c) ------------------------------------------------------------------
    def closed_loop(self, model, N, path, pose0, ticks, act0=None, v=1.0, T=0.1, l=0.21, lookahead=None, P=64,
                    eps=0.01, max_iter=10000, nthreads=8, horizon=None):
        """horizon = (d0, step): a target per horizon step sampled at d_t = d0 + t*step (orc_closed_loop_h)."""
        ma = _ModelArgs(model)
        px, py, tx, ty = (_c64(a) for a in path)
        pose = _c64(pose0).copy(); n = pose.shape[0]
        act = np.zeros((n, 2)) if act0 is None else _c64(act0).copy()
        u0 = np.zeros(n); tgt = np.zeros((n, 2)); idx = np.zeros(n, dtype=np.int32)
        itot = np.zeros(n, dtype=np.uint64); imax = np.zeros(n, dtype=np.uint32)
        f = self.lib.orc_closed_loop_h
        f.restype = C.c_int
        f.argtypes = [C.c_long, C.c_int] + [_dp] * 7 + [C.c_double, C.c_ulong] + [C.c_double] * 4 + [C.c_int, C.c_int] + \
            [_dp] * 4 + [C.c_int] + [_dp] * 4 + [C.POINTER(C.c_int32), C.POINTER(C.c_uint64), C.POINTER(C.c_uint32), C.c_int] + \
            [C.c_int, C.c_float, C.c_float]
        hz = (0, 0.0, 0.0) if horizon is None else (1, float(np.float32(horizon[0])), float(np.float32(horizon[1])))
        rc = f(n, N, *ma.ptrs, eps, max_iter, v, T, l, (0.6 + 0.1 * v) if lookahead is None else lookahead, P,
               px.shape[0], _d(px), _d(py), _d(tx), _d(ty), ticks, _d(pose), _d(act), _d(u0), _d(tgt),
               idx.ctypes.data_as(C.POINTER(C.c_int32)), itot.ctypes.data_as(C.POINTER(C.c_uint64)),
               imax.ctypes.data_as(C.POINTER(C.c_uint32)), nthreads, *hz)
        if rc != 0:
            raise ValueError("oracle: bad closed-loop arguments")
        return dict(pose=pose, act=act, u0=u0, target=tgt, idx=idx, iters_total=itot, iters_max=imax)

    # ---- look-up table, crossing clamp, PID branch (oracle/pid_oracle.c) ------------------------------------------
    def lookup(self, vx, vy, x):
        vx = np.ascontiguousarray(vx, dtype=np.float32); vy = np.ascontiguousarray(vy, dtype=np.float32)
        f = self.lib.orc_lookup
        f.restype = C.c_float
        f.argtypes = [_fp, _fp, C.c_int, C.c_float]
        return np.array([f(_f(vx), _f(vy), len(vx), float(v)) for v in np.asarray(x, dtype=np.float32)], dtype=np.float32)

    def crossing(self, px, py, vel, point_velocity, state, gains=(1.0, 0.0, 0.0, 0.01), cfg=(1.0, 0.35, 1.0, 0.05),
                 npoints=None, layout="pb"):
        """src/trajectory_point_follower.cpp:445-473 for a batch; `state` = (integral, prev, first) arrays, updated in
        place like `point_velocity`."""
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, vel)]
        B = point_velocity.shape[0]
        if layout == "shared":
            P, si, sp = arrs[0].shape[0], 0, 1
        elif layout == "pb":
            P, si, sp = arrs[0].shape[0], 1, B
        else:
            P, si, sp = arrs[0].shape[1], arrs[0].shape[1], 1
        npts = None if npoints is None else np.ascontiguousarray(npoints, dtype=np.int32)
        c = (C.c_float * 4)(*cfg)
        f = self.lib.orc_crossing_batch
        f.restype = None
        f.argtypes = [C.c_long, C.c_int, C.POINTER(C.c_int32), _fp, _fp, _fp, C.c_long, C.c_long, C.c_void_p] + \
            [C.c_float] * 4 + [_fp, _fp, C.POINTER(C.c_uint8), _fp]
        f(B, P, None if npts is None else npts.ctypes.data_as(C.POINTER(C.c_int32)), *[_f(a) for a in arrs], si, sp,
          C.cast(c, C.c_void_p), *[float(g) for g in gains], _f(state[0]), _f(state[1]),
          state[2].ctypes.data_as(C.POINTER(C.c_uint8)), _f(point_velocity))

    def pid_branch(self, pos_x, pos_y, dir_x, dir_y, gains_front, gains_rear, state_front, state_rear):
        B = pos_x.shape[0]
        sf = np.zeros(B); sr = np.zeros(B); td = np.zeros(B, dtype=np.float32)
        gf = (C.c_float * 4)(*gains_front); gr = (C.c_float * 4)(*gains_rear)
        u8 = C.POINTER(C.c_uint8)
        f = self.lib.orc_pid_branch_batch
        f.restype = None
        f.argtypes = [C.c_long] + [_fp] * 4 + [C.c_void_p, C.c_void_p, _fp, _fp, u8, _fp, _fp, u8, _dp, _dp, _fp]
        f(B, _f(pos_x), _f(pos_y), _f(dir_x), _f(dir_y), C.cast(gf, C.c_void_p), C.cast(gr, C.c_void_p),
          _f(state_front[0]), _f(state_front[1]), state_front[2].ctypes.data_as(u8),
          _f(state_rear[0]), _f(state_rear[1]), state_rear[2].ctypes.data_as(u8), _d(sf), _d(sr), _f(td))
        return sf, sr, td

    # ---- point extraction ---------------------------------------------------------------------------------
    def extract_horizon(self, px, py, dx, dy, vel, d0, step, steps, npoints=None, layout="shared"):
        """orc_extract_horizon per instance: (idx[B][steps], flag[B][steps], (px, py, dx, dy, vel)[B][steps])."""
        class Pt(C.Structure):
            _fields_ = [("pos_x", C.c_float), ("pos_y", C.c_float), ("dir_x", C.c_float), ("dir_y", C.c_float), ("vel", C.c_float),
                        ("idx", C.c_int32), ("flag", C.c_uint8)]
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, dx, dy, vel)]
        d0 = np.ascontiguousarray(d0, dtype=np.float32); step = np.ascontiguousarray(step, dtype=np.float32)
        B = d0.shape[0]
        if layout == "shared":
            P, si, sp = arrs[0].shape[0], 0, 1
        elif layout == "pb":
            P, si, sp = arrs[0].shape[0], 1, B
        else:
            P, si, sp = arrs[0].shape[1], arrs[0].shape[1], 1
        f = self.lib.orc_extract_horizon
        f.restype = None
        f.argtypes = [C.c_void_p] * 5 + [C.c_long, C.c_int, C.c_float, C.c_float, C.c_int, C.POINTER(Pt)]
        idx = np.zeros((B, steps), dtype=np.int32); flag = np.zeros((B, steps), dtype=np.uint8)
        outs = [np.zeros((B, steps), dtype=np.float32) for _ in range(5)]
        pts = (Pt * steps)()
        for b in range(B):
            n = P if npoints is None else min(int(npoints[b]), P)
            f(*[a.ctypes.data + 4 * b * si for a in arrs], sp, n, float(d0[b]), float(step[b]), steps, pts)
            for t in range(steps):
                idx[b, t] = pts[t].idx; flag[b, t] = pts[t].flag
                for o, k in zip(outs, ("pos_x", "pos_y", "dir_x", "dir_y", "vel")):
                    o[b, t] = getattr(pts[t], k)
        return idx, flag, tuple(outs)

    def extract(self, px, py, dx, dy, vel, dist, npoints=None, layout="shared"):
        """layout 'shared': component arrays [P]; 'pb': [P][B] (point-major); 'bp': [B][P]."""
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, dx, dy, vel)]
        dist = np.ascontiguousarray(dist, dtype=np.float32)
        B = dist.shape[0]
        if layout == "shared":
            P, si, sp = arrs[0].shape[0], 0, 1
        elif layout == "pb":
            P, si, sp = arrs[0].shape[0], 1, B
        else:
            P, si, sp = arrs[0].shape[1], arrs[0].shape[1], 1
        npts = None if npoints is None else np.ascontiguousarray(npoints, dtype=np.int32)
        idx = np.zeros(B, dtype=np.int32); flag = np.zeros(B, dtype=np.uint8)
        outs = [np.zeros(B, dtype=np.float32) for _ in range(5)]
        self.lib.orc_extract_batch(B, P, None if npts is None else npts.ctypes.data_as(C.POINTER(C.c_int32)),
                                   *[_f(a) for a in arrs], si, sp, _f(dist),
                                   idx.ctypes.data_as(C.POINTER(C.c_int32)),
                                   flag.ctypes.data_as(C.POINTER(C.c_uint8)), *[_f(o) for o in outs])
        return idx, flag, outs


class _OrcStruct(C.Structure):
    MAX_S, MAX_I, MAX_N = 8, 4, 128
    _fields_ = [("S", C.c_int), ("I", C.c_int), ("N", C.c_int),
                ("A", C.c_double * 64), ("B", C.c_double * 32), ("C", C.c_double * 8), ("Q", C.c_double * 8),
                ("R", C.c_double * 4), ("lower", C.c_double * 4), ("upper", C.c_double * 4),
                ("max_iterations", C.c_ulong), ("eps", C.c_double), ("lam", C.c_double),
                ("Q_diag", C.c_double * 512), ("target", C.c_double * 1024), ("controls", C.c_double * 512),
                ("v", C.c_double * 512), ("last_iters", C.c_ulong), ("last_max_df", C.c_double)]


class OracleController:
    """One persistent orc_mpc (mirrors a dlib::mpc object)."""

    def __init__(self, orc: Oracle, model, N: int):
        self.orc, self.N = orc, N
        ma = _ModelArgs(model)
        self.S, self.I = ma.S, ma.I
        self.st = _OrcStruct()
        rc = orc.lib.orc_mpc_init(C.byref(self.st), ma.S, ma.I, N, *ma.ptrs)
        if rc != 0:
            raise ValueError("oracle: shape out of bounds")

    def set_options(self, eps=None, max_iter=None):
        if eps is not None:
            self.st.eps = eps
        if max_iter is not None:
            self.st.max_iterations = max_iter

    def set_target(self, val, time=-1):
        v = _c64(val)
        self.orc.lib.orc_mpc_set_target(C.byref(self.st), _d(v), time)

    def step(self, x0):
        x = _c64(x0); u = np.zeros(self.I)
        self.orc.lib.orc_mpc_step(C.byref(self.st), _d(x), _d(u))
        return u

    @property
    def iters(self):
        return int(self.st.last_iters)

    @property
    def lam(self):
        return float(self.st.lam)

    def q_diag(self):
        return np.array(self.st.Q_diag[: self.N * self.I]).reshape(self.N, self.I)

    def controls(self):
        return np.array(self.st.controls[: self.N * self.I]).reshape(self.N, self.I)

    def v(self):
        return np.array(self.st.v[: self.N * self.I]).reshape(self.N, self.I)

    def targets(self):
        return np.array(self.st.target[: self.N * self.S]).reshape(self.N, self.S)


class DlibRef:
    """The real dlib::mpc (oracle/_ref/libdlibref.so).  ``DlibRef.available()`` is False on the GPU box unless
    the prebuilt library travelled with the snapshot."""
    PATH = os.path.join(ORACLE_DIR, "_ref", "libdlibref.so")

    @classmethod
    def available(cls) -> bool:
        return os.path.exists(cls.PATH)

    def __init__(self):
        self.lib = L = C.CDLL(self.PATH)
        L.dlibref_supported.argtypes = [C.c_int] * 3
        L.dlibref_new.restype = C.c_void_p
        L.dlibref_new.argtypes = [C.c_int] * 3 + [_dp] * 7
        L.dlibref_free.argtypes = [C.c_void_p]
        L.dlibref_set_target.argtypes = [C.c_void_p, _dp, C.c_long]
        L.dlibref_set_options.argtypes = [C.c_void_p, C.c_double, C.c_ulong]
        L.dlibref_step.argtypes = [C.c_void_p, _dp, _dp]
        L.dlibref_get_state.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.dlibref_set_state.argtypes = [C.c_void_p, _dp, _dp, _dp]
        L.dlibref_get_derived.argtypes = [C.c_void_p, _dp, _dp]
        L.dlibref_batch_cold.restype = C.c_double
        L.dlibref_batch_cold.argtypes = [C.c_int] * 4 + [_dp] * 7 + [C.c_long, _dp, _dp, C.c_double, C.c_ulong,
                                                                    _dp, _dp, C.c_int]

    def supported(self, S, I, N) -> bool:
        return bool(self.lib.dlibref_supported(S, I, N))

    def batch_cold(self, model, N, x0, target=None, eps=0.01, max_iter=10000, per_instance=False, nthreads=1,
                   want_controls=True):
        ma = _ModelArgs(model, per_instance)
        x0 = _c64(x0); B = x0.shape[0]
        tgt = None if target is None else _c64(target)
        u0 = np.zeros((B, ma.I)); ctr = np.zeros((B, N, ma.I)) if want_controls else None
        secs = self.lib.dlibref_batch_cold(ma.S, ma.I, N, int(per_instance), *ma.ptrs, B, _d(x0), _d(tgt), eps,
                                           max_iter, _d(u0), _d(ctr), nthreads)
        if secs < 0:
            raise ValueError(f"dlibref: shape ({ma.S},{ma.I},{N}) not instantiated")
        return u0, ctr, secs

    def controller(self, model, N):
        return DlibController(self, model, N)


class DlibController:
    def __init__(self, ref: DlibRef, model, N: int):
        ma = _ModelArgs(model)
        self.ref, self.N, self.S, self.I = ref, N, ma.S, ma.I
        self.h = ref.lib.dlibref_new(ma.S, ma.I, N, *ma.ptrs)
        if not self.h:
            raise ValueError(f"dlibref: shape ({ma.S},{ma.I},{N}) not instantiated")

    def __del__(self):
        if getattr(self, "h", None):
            self.ref.lib.dlibref_free(self.h)
            self.h = None

    def set_options(self, eps, max_iter):
        self.ref.lib.dlibref_set_options(self.h, eps, max_iter)

    def set_target(self, val, time=-1):
        v = _c64(val)
        self.ref.lib.dlibref_set_target(self.h, _d(v), time)

    def step(self, x0):
        x = _c64(x0); u = np.zeros(self.I)
        self.ref.lib.dlibref_step(self.h, _d(x), _d(u))
        return u

    def controls(self):
        c = np.zeros((self.N, self.I)); self.ref.lib.dlibref_get_state(self.h, _d(c), None, None); return c

    def v(self):
        c = np.zeros((self.N, self.I)); self.ref.lib.dlibref_get_state(self.h, None, None, _d(c)); return c

    def targets(self):
        c = np.zeros((self.N, self.S)); self.ref.lib.dlibref_get_state(self.h, None, _d(c), None); return c

    def derived(self):
        lam = C.c_double(); qd = np.zeros((self.N, self.I))
        self.ref.lib.dlibref_get_derived(self.h, C.byref(lam), _d(qd))
        return lam.value, qd


def bits(a) -> np.ndarray:
    """uint64 bit patterns of a float64 array (for bit-exact comparisons)."""
    return np.ascontiguousarray(a, dtype=np.float64).view(np.uint64)


def same_bits(a, b) -> bool:
    return bool(np.array_equal(bits(a), bits(b)))


def same_values(a, b) -> bool:
    """IEEE equality that also accepts -0 == +0 and NaN == NaN: the only freedom the CUDA path takes
    (it does not multiply by diagm()'s literal zeros, which can only flip the sign of a zero)."""
    a = np.asarray(a, dtype=np.float64); b = np.asarray(b, dtype=np.float64)
    return bool(np.all((a == b) | (np.isnan(a) & np.isnan(b))))


def fnv1a64_of_doubles(vals) -> str:
    """FNV-1a-64 over the little-endian bytes of the doubles (the golden hash format of BASELINE.md §2)."""
    h = 0xCBF29CE484222325
    for byte in np.ascontiguousarray(vals, dtype="<f8").tobytes():
        h = ((h ^ byte) * 0x100000001B3) & 0xFFFFFFFFFFFFFFFF
    return f"{h:016x}"
