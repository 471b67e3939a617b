"""ctypes binding of the C-ABI product library ``libbmpc.so`` (include/bmpc.h).

This is what a Python caller of the engine uses; tests and bench.py go through it, i.e. through the C ABI.
There is no fallback of any kind: if the library is missing, not loadable, or no CUDA device is present, the
calls raise.  Nothing in here imports or touches ``oracle/``.

``BatchedMPC`` mirrors ``dlib::mpc<S,I,N>``'s method set (/root/reference/dlib_files/dlib/control/mpc.h:127-240)
for a batch of controllers: ``set_target``, ``set_last_target``, ``get_target``, ``set_epsilon``,
``set_max_iterations``, and ``__call__(x0)`` (= operator()) with the persistent-controller shift semantics.
"""
from __future__ import annotations

import ctypes as C
import os

import numpy as np

_HERE = os.path.dirname(os.path.abspath(__file__))
# BMPC_LIB: developer override to A/B-test another build of the same ABI (never a fallback: the file must exist)
LIB_PATH = os.environ.get("BMPC_LIB") or os.path.join(_HERE, "libbmpc.so")

FP64, FP32 = 64, 32
EXACT, FAST = 0, 1
WARM, KEEP_CONTROLS = 1, 2
ST_MAXITER = 1
ST_NONFINITE = 2
PT_INTERPOLATED, PT_LAST, PT_EMPTY = 0, 1, 2

_dp = C.POINTER(C.c_double)
_fp = C.POINTER(C.c_float)
_i32p = C.POINTER(C.c_int32)
_u32p = C.POINTER(C.c_uint32)
_u8p = C.POINTER(C.c_uint8)

# name -> (restype, argtypes): every symbol include/bmpc.h declares
SIGNATURES = {
    "bmpc_shape_supported": (C.c_int, [C.c_int] * 3),
    "bmpc_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int, C.c_int]),
    "bmpc_destroy": (None, [C.c_void_p]),
    "bmpc_last_error": (C.c_char_p, []),
    "bmpc_version": (C.c_char_p, []),
    "bmpc_set_model": (C.c_int, [C.c_void_p] + [_dp] * 7 + [C.c_int, C.c_int64]),
    "bmpc_set_specialization": (C.c_int, [C.c_void_p, C.c_int]),
    "bmpc_set_split": (C.c_int, [C.c_void_p, C.c_int]),
    "bmpc_pattern_name": (C.c_char_p, [C.c_void_p]),
    "bmpc_set_options": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64]),
    "bmpc_get_options": (C.c_int, [C.c_void_p, _dp, C.POINTER(C.c_uint64)]),
    "bmpc_get_derived": (C.c_int, [C.c_void_p, C.c_int64, _dp, _dp]),
    "bmpc_set_target": (C.c_int, [C.c_void_p, _dp, C.c_int, C.c_int, C.c_int64]),
    "bmpc_get_target": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, _dp]),
    "bmpc_reset": (C.c_int, [C.c_void_p]),
    "bmpc_get_controls": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, _dp]),
    "bmpc_set_controls": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, _dp]),
    "bmpc_copy_state": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bmpc_solve": (C.c_int, [C.c_void_p, C.c_int64, _dp, _dp, _u32p, _u8p, C.c_uint]),
    "bmpc_upload_x0": (C.c_int, [C.c_void_p, C.c_int64, _dp]),
    "bmpc_solve_resident_async": (C.c_int, [C.c_void_p, C.c_int64, C.c_uint]),
    "bmpc_download": (C.c_int, [C.c_void_p, C.c_int64, _dp, _u32p, _u8p]),
    "bmpc_sync": (C.c_int, [C.c_void_p]),
    "bmpc_device_results": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_void_p)] * 3),
    "bmpc_device_results_block": (C.c_int, [C.c_void_p, C.POINTER(C.c_void_p)] + [C.POINTER(C.c_uint64)] * 3),
    "bmpc_set_stream": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bmpc_last_kernel_ms": (C.c_int, [C.c_void_p, _fp]),
    "bmpc_last_stage_ms": (C.c_int, [C.c_void_p, _fp, _fp]),
    "bmpc_pg_kernel_info": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_int)] * 4),
    "bmpc_launch_count": (C.c_int64, [C.c_void_p]),
    "bmpc_kernel_info": (C.c_int, [C.c_void_p] + [C.POINTER(C.c_int)] * 4),
    "bmpc_probe_peak": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _dp, _dp]),
    "bmpc_extract_points": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, _i32p] + [_fp] * 5 + [C.c_int64, C.c_int64, _fp, _i32p, _u8p] + [_fp] * 5),
    "bmpc_extract_points_static": (C.c_int, [C.c_int, C.c_void_p, C.c_int64, C.c_int, _i32p] + [_fp] * 5 +
                                   [C.c_int64, C.c_int64, _fp, _i32p, _u8p] + [_fp] * 5),
    "bmpc_fleet_tick_tobi": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p, _fp, _dp, _fp, C.c_int, _i32p] + [_fp] * 5 +
                             [C.c_int64, C.c_int64, _dp, _dp, _fp, _fp, _u8p, _i32p, _u8p, _dp, _u32p]),
    "bmpc_set_lookup": (C.c_int, [C.c_void_p, C.c_int, C.c_int, _fp, _fp]),
    "bmpc_lookup": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _fp, _fp]),
    "bmpc_fleet_set_drive_mode": (C.c_int, [C.c_void_p, C.c_int, C.c_float, C.c_float]),
    "bmpc_fleet_set_crossing": (C.c_int, [C.c_void_p, C.c_void_p]),
    "bmpc_fleet_set_pid": (C.c_int, [C.c_void_p, C.c_int] + [C.c_float] * 4),
    "bmpc_fleet_reset_pid": (C.c_int, [C.c_void_p, C.c_int]),
    "bmpc_fleet_get_pid": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _fp, _fp, _u8p]),
    "bmpc_fleet_put_pid": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, _fp, _fp, _u8p]),
    "bmpc_fleet_tick_pid": (C.c_int, [C.c_void_p, C.c_int64, _fp, _fp, C.c_int, _i32p] + [_fp] * 5 +
                            [C.c_int64, C.c_int64, _dp, _dp, _fp, _fp, _i32p, _u8p]),
    "bmpc_cl_init": (C.c_int, [C.c_void_p, C.c_int64, C.c_void_p] + [_dp] * 6),
    "bmpc_cl_run": (C.c_int, [C.c_void_p, C.c_int]),
    "bmpc_cl_set_horizon_targets": (C.c_int, [C.c_void_p, C.c_int, C.c_double, C.c_double]),
    "bmpc_extract_horizon": (C.c_int, [C.c_void_p, C.c_int64, C.c_int, _i32p] + [_fp] * 5 + [C.c_int64, C.c_int64, _fp, _fp, C.c_int,
                                       _i32p, _u8p] + [_fp] * 5),
    "bmpc_cl_get": (C.c_int, [C.c_void_p, C.c_int64, _dp, _dp, _dp, _dp, _i32p, C.POINTER(C.c_uint64), _u32p]),
    "bmpc_group_create": (C.c_int, [C.POINTER(C.c_void_p), C.c_int, C.c_int, C.c_int, C.c_int64, C.c_int, C.c_int,
                                    C.POINTER(C.c_int), C.c_int, C.c_int]),
    "bmpc_group_destroy": (None, [C.c_void_p]),
    "bmpc_group_size": (C.c_int, [C.c_void_p]),
    "bmpc_group_member": (C.c_void_p, [C.c_void_p, C.c_int]),
    "bmpc_group_shard": (C.c_int, [C.c_void_p, C.c_int, C.c_int64, C.POINTER(C.c_int64), C.POINTER(C.c_int64)]),
    "bmpc_group_set_model": (C.c_int, [C.c_void_p] + [_dp] * 7 + [C.c_int, C.c_int64]),
    "bmpc_group_set_options": (C.c_int, [C.c_void_p, C.c_double, C.c_uint64]),
    "bmpc_group_set_target": (C.c_int, [C.c_void_p, _dp, C.c_int, C.c_int, C.c_int64]),
    "bmpc_group_reset": (C.c_int, [C.c_void_p]),
    "bmpc_group_solve": (C.c_int, [C.c_void_p, C.c_int64, _dp, _dp, _u32p, _u8p, C.c_uint]),
    "bmpc_group_get_controls": (C.c_int, [C.c_void_p, C.c_int64, C.c_int64, _dp]),
    "bmpc_group_last_kernel_ms": (C.c_int, [C.c_void_p, _fp]),
}


class ClConfig(C.Structure):
    """bmpc_cl_config (include/bmpc.h)."""
    _fields_ = [("v", C.c_double), ("T", C.c_double), ("l", C.c_double), ("lookahead", C.c_double),
                ("P", C.c_int32), ("npath", C.c_int32)]


_lib = None


class TickConfig(C.Structure):
    """bmpc_tick_config; defaults = the reference's (src/trajectory_point_follower.cpp:16,92-96, header :47)."""
    _fields_ = [("T", C.c_double), ("l", C.c_double), ("weight_y", C.c_double), ("weight_phi", C.c_double),
                ("weight_steering_front", C.c_double), ("weight_steering_rear", C.c_double), ("alpha_max", C.c_double)]

    def __init__(self, T=0.1, l=0.21, weight_y=20.0, weight_phi=7.0, weight_steering_front=0.0005, weight_steering_rear=10.0,
                 alpha_max=22 * 3.141592653589793 / 180):
        super().__init__(T, l, weight_y, weight_phi, weight_steering_front, weight_steering_rear, alpha_max)


class CrossingConfig(C.Structure):
    """bmpc_crossing_config; defaults = the reference's config defaults (src/trajectory_point_follower.cpp:446-457)."""
    _fields_ = [("max_velocity_crossing", C.c_float), ("stopping_distance", C.c_float), ("distance_to_stop", C.c_float),
                ("crossing_safety_zone", C.c_float)]

    def __init__(self, max_velocity_crossing=1.0, stopping_distance=0.35, distance_to_stop=1.0, crossing_safety_zone=0.05):
        super().__init__(max_velocity_crossing, stopping_distance, distance_to_stop, crossing_safety_zone)


LOOKUP_DISTANCE, LOOKUP_VELOCITY = 0, 1
GROUP_INTERLEAVED, GROUP_CONTIGUOUS = 0, 1
PID_SLOWDOWN, PID_FRONT, PID_REAR = 0, 1, 2


class BmpcError(RuntimeError):
    def __init__(self, code: int, msg: str):
        super().__init__(f"bmpc error {code}: {msg}")
        self.code = code


def lib() -> C.CDLL:
    """Load libbmpc.so (once).  Raises if it has not been built — there is no fallback path."""
    global _lib
    if _lib is None:
        if not os.path.exists(LIB_PATH):
            raise FileNotFoundError(f"{LIB_PATH} not built: run `python -c 'import __graft_entry__ as g; g.build()'` "
                                    "or `make -C trajectory_controller_b200/csrc -j8`")
        L = C.CDLL(LIB_PATH)
        for name, (res, args) in SIGNATURES.items():
            if os.environ.get("BMPC_LIB") and not hasattr(L, name):
                continue                      # developer A/B builds of older sources (scripts/): newer entry points absent
            fn = getattr(L, name)
            fn.restype, fn.argtypes = res, args
        _lib = L
    return _lib


def _check(rc: int):
    if rc != 0:
        raise BmpcError(rc, lib().bmpc_last_error().decode())


def _c64(a):
    return np.ascontiguousarray(a, dtype=np.float64)


def _p(a, t=_dp):
    return None if a is None else a.ctypes.data_as(t)


def shape_supported(S: int, I: int, N: int) -> bool:
    return bool(lib().bmpc_shape_supported(S, I, N))


class _DeviceView:
    """Minimal CUDA array interface wrapper around a device pointer owned by a handle."""

    def __init__(self, ptr: int, shape, typestr: str):
        self.__cuda_array_interface__ = {"shape": tuple(shape), "typestr": typestr, "data": (int(ptr), False),
                                         "version": 2, "strides": None}


class BatchedMPC:
    """`capacity` dlib::mpc<S,I,N>-compatible controllers on one GPU."""

    def __init__(self, S: int, I: int, N: int, capacity: int, precision: int = FP64, mode: int = EXACT, device: int = -1):
        self.S, self.I, self.N, self.capacity = S, I, N, int(capacity)
        self.precision, self.mode = precision, mode
        self._h = C.c_void_p()
        _check(lib().bmpc_create(C.byref(self._h), S, I, N, self.capacity, precision, mode, device))
        if os.environ.get("BMPC_SPLIT"):   # developer A/B knob (scripts/): the bmpc_set_split setting of every handle
            self.set_split(int(os.environ["BMPC_SPLIT"]))

    def close(self):
        if getattr(self, "_h", None) is not None and self._h:
            lib().bmpc_destroy(self._h)
            self._h = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    # ---- model / options -----------------------------------------------------------------------------------------
    def set_model(self, model, per_instance: bool = False):
        g = (lambda k: model[k]) if isinstance(model, dict) else (lambda k: getattr(model, k))
        arrs = [_c64(g(k)) for k in ("A", "B", "C", "Q", "R", "lower", "upper")]
        count = arrs[0].shape[0] if per_instance else 1
        _check(lib().bmpc_set_model(self._h, *[_p(a) for a in arrs], int(per_instance), count))
        return self

    def set_specialization(self, enable: bool):
        """False: always run the dense kernel (effective at the next set_model)."""
        _check(lib().bmpc_set_specialization(self._h, int(enable)))
        return self

    def set_split(self, enable):
        """0/False: one-kernel solve; 1/True: two-stage solve (coordinate-phase kernel + projected-gradient kernel)
        with fairness slices of small batches, retirement of surplus warps and the symmetric-bound stage-B kernel
        (default); 2: two-stage, every instance runs to completion in the lane that claimed it; 3: like 1 with the
        literal stage-B kernel even where the bounds are symmetric.  Results are identical."""
        _check(lib().bmpc_set_split(self._h, int(enable)))

    def pattern_name(self) -> str:
        return lib().bmpc_pattern_name(self._h).decode()

    def set_epsilon(self, eps: float):
        _check(lib().bmpc_set_options(self._h, eps, self.get_max_iterations()))

    def set_max_iterations(self, n: int):
        _check(lib().bmpc_set_options(self._h, self.get_epsilon(), n))

    def set_options(self, eps: float, max_iter: int):
        _check(lib().bmpc_set_options(self._h, eps, max_iter))

    def _opts(self):
        e, m = C.c_double(), C.c_uint64()
        _check(lib().bmpc_get_options(self._h, C.byref(e), C.byref(m)))
        return e.value, m.value

    def get_epsilon(self) -> float:
        return self._opts()[0]

    def get_max_iterations(self) -> int:
        return self._opts()[1]

    def derived(self, instance: int = 0):
        lam = C.c_double(); qd = np.zeros((self.N, self.I))
        _check(lib().bmpc_get_derived(self._h, instance, C.byref(lam), _p(qd)))
        return lam.value, qd

    # ---- targets ---------------------------------------------------------------------------------------------------
    def set_target(self, val, time: int = -1):
        """val: [S] (all instances) or [count][S].  time=-1: every horizon step (mpc.h:157-163)."""
        v = _c64(val)
        per = v.ndim == 2
        _check(lib().bmpc_set_target(self._h, _p(v), time, int(per), v.shape[0] if per else 0))

    def set_last_target(self, val):
        self.set_target(val, self.N - 1)

    def get_target(self, time: int, instance: int = 0):
        out = np.zeros(self.S)
        _check(lib().bmpc_get_target(self._h, instance, time, _p(out)))
        return out

    # ---- state --------------------------------------------------------------------------------------------------------
    def reset(self):
        _check(lib().bmpc_reset(self._h))

    def get_controls(self, count: int | None = None, first: int = 0):
        count = self.capacity - first if count is None else count
        out = np.zeros((count, self.N, self.I))
        _check(lib().bmpc_get_controls(self._h, first, count, _p(out)))
        return out

    def set_controls(self, controls, first: int = 0):
        c = _c64(controls)
        _check(lib().bmpc_set_controls(self._h, first, c.shape[0], _p(c)))

    # ---- solve ----------------------------------------------------------------------------------------------------------
    def solve(self, x0, warm: bool = False, keep_controls: bool = False, out=None):
        """operator() for x0[count][S].  Returns (u0[count][I], iters[count] uint32, status[count] uint8)."""
        x = _c64(x0)
        B = x.shape[0]
        if out is None:
            u0 = np.empty((B, self.I)); it = np.empty(B, dtype=np.uint32); st = np.empty(B, dtype=np.uint8)
        else:
            u0, it, st = out
        flags = (WARM if warm else 0) | (KEEP_CONTROLS if keep_controls else 0)
        _check(lib().bmpc_solve(self._h, B, _p(x), _p(u0), _p(it, _u32p), _p(st, _u8p), flags))
        return u0, it, st

    __call__ = solve

    def upload_x0(self, x0):
        x = _c64(x0)
        _check(lib().bmpc_upload_x0(self._h, x.shape[0], _p(x)))
        return x.shape[0]

    def solve_resident_async(self, count: int, warm: bool = False, keep_controls: bool = False):
        _check(lib().bmpc_solve_resident_async(self._h, count, (WARM if warm else 0) | (KEEP_CONTROLS if keep_controls else 0)))

    def download(self, count: int, out=None):
        if out is None:
            u0 = np.empty((count, self.I)); it = np.empty(count, dtype=np.uint32); st = np.empty(count, dtype=np.uint8)
        else:
            u0, it, st = out
        _check(lib().bmpc_download(self._h, count, _p(u0), _p(it, _u32p), _p(st, _u8p)))
        return u0, it, st

    def sync(self):
        _check(lib().bmpc_sync(self._h))

    def device_results(self):
        """Device views (``__cuda_array_interface__`` objects, e.g. for ``torch.as_tensor(v, device="cuda")``) of the last
        solve's results in HBM: u0 [I][capacity] in the compute type, iters uint32 (exposed as int32) [capacity],
        status uint8 [capacity].  For callers that gather results with a collective instead of a D2H per GPU."""
        ptrs = [C.c_void_p() for _ in range(3)]
        _check(lib().bmpc_device_results(self._h, *[C.byref(p) for p in ptrs]))
        ft = "<f8" if self.precision == FP64 else "<f4"
        return (_DeviceView(ptrs[0].value, (self.I, self.capacity), ft),
                _DeviceView(ptrs[1].value, (self.capacity,), "<i4"),
                _DeviceView(ptrs[2].value, (self.capacity,), "|u1"))

    def device_results_block(self):
        """(device view of the whole result block as uint8 [bytes], iters_offset, status_offset): u0 | iters | status in
        one allocation — one collective gathers a rank's results."""
        base = C.c_void_p(); n, oi, os_ = C.c_uint64(), C.c_uint64(), C.c_uint64()
        _check(lib().bmpc_device_results_block(self._h, C.byref(base), C.byref(n), C.byref(oi), C.byref(os_)))
        return _DeviceView(base.value, (n.value,), "|u1"), oi.value, os_.value

    def set_stream(self, cuda_stream: int | None):
        """Borrow a caller's cudaStream_t (integer handle, e.g. torch.cuda.current_stream().cuda_stream); None/0 = own."""
        _check(lib().bmpc_set_stream(self._h, C.c_void_p(cuda_stream or 0)))

    def last_kernel_ms(self) -> float:
        ms = C.c_float()
        _check(lib().bmpc_last_kernel_ms(self._h, C.byref(ms)))
        return ms.value

    def last_stage_ms(self):
        a, b = C.c_float(), C.c_float()
        _check(lib().bmpc_last_stage_ms(self._h, C.byref(a), C.byref(b)))
        return a.value, b.value

    def pg_kernel_info(self):
        v = [C.c_int() for _ in range(4)]
        _check(lib().bmpc_pg_kernel_info(self._h, *[C.byref(x) for x in v]))
        return dict(zip(("grid", "block", "smem_bytes", "regs"), (x.value for x in v)))

    def launch_count(self) -> int:
        return int(lib().bmpc_launch_count(self._h))

    def kernel_info(self):
        g, b, s, r = C.c_int(), C.c_int(), C.c_int(), C.c_int()
        _check(lib().bmpc_kernel_info(self._h, C.byref(g), C.byref(b), C.byref(s), C.byref(r)))
        return dict(grid=g.value, block=b.value, smem_bytes=s.value, regs=r.value)

    def probe_peak(self, precision: int = FP64, fused: bool = True):
        """(TFLOP/s, G instructions/s) of the device's FP64/FP32 pipe, measured with register-resident chains."""
        t, g = C.c_double(), C.c_double()
        _check(lib().bmpc_probe_peak(self._h, precision, int(fused), C.byref(t), C.byref(g)))
        return t.value, g.value

    # ---- closed loop on the device ------------------------------------------------------------------------------------------
    def cl_init(self, path, pose0, act0=None, v=1.0, T=0.1, l=0.21, lookahead=None, P=64):
        """path: (x, y, tx, ty) world polyline arrays; pose0 [count][4] = X, Y, cos psi, sin psi."""
        px, py, tx, ty = (_c64(a) for a in path)
        pose0 = _c64(pose0)
        cfg = ClConfig(v, T, l, (0.6 + 0.1 * v) if lookahead is None else lookahead, P, px.shape[0])
        a0 = None if act0 is None else _c64(act0)
        _check(lib().bmpc_cl_init(self._h, pose0.shape[0], C.byref(cfg), _p(px), _p(py), _p(tx), _p(ty), _p(pose0), _p(a0)))
        self._cl_count = pose0.shape[0]

    def cl_set_horizon_targets(self, enable: bool = True, d0: float | None = None, step: float | None = None, v=1.0, T=0.1):
        """A target per horizon step, sampled at d_t = d0 + t*step (defaults d0 = step = T*v); see include/bmpc.h."""
        d0 = T * v if d0 is None else d0
        step = T * v if step is None else step
        _check(lib().bmpc_cl_set_horizon_targets(self._h, int(enable), d0, step))

    def cl_run(self, ticks: int):
        _check(lib().bmpc_cl_run(self._h, ticks))

    def cl_get(self):
        n = self._cl_count
        out = dict(pose=np.empty((n, 4)), act=np.empty((n, 2)), u0=np.empty(n), target=np.empty((n, 2)),
                   idx=np.empty(n, dtype=np.int32), iters_total=np.empty(n, dtype=np.uint64), iters_max=np.empty(n, dtype=np.uint32))
        _check(lib().bmpc_cl_get(self._h, n, _p(out["pose"]), _p(out["act"]), _p(out["u0"]), _p(out["target"]),
                                 _p(out["idx"], _i32p), _p(out["iters_total"], C.POINTER(C.c_uint64)), _p(out["iters_max"], _u32p)))
        return out

    # ---- fleet tick (TrajectoryPointController::cycle, tobiMPC branch, batched) ------------------------------------------
    def fleet_tick_tobi(self, velocity, dist, px, py, dx, dy, vel, layout: str = "pb", v_model=None, npoints=None, cfg=None):
        """One control tick for len(velocity) vehicles: returns a dict of CarCommand fields and diagnostics."""
        cfg = cfg or TickConfig()
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, dx, dy, vel)]
        velocity = np.ascontiguousarray(velocity, dtype=np.float32)
        dist = None if dist is None else np.ascontiguousarray(dist, dtype=np.float32)   # None: look-ahead on the device
        B = velocity.shape[0]
        P, si, sp = self._poly_strides(arrs[0], layout, B)
        vm = None if v_model is None else _c64(v_model)
        npts = None if npoints is None else np.ascontiguousarray(npoints, dtype=np.int32)
        out = dict(steering_front=np.empty(B), steering_rear=np.empty(B), target_speed=np.empty(B, dtype=np.float32),
                   target_distance=np.empty(B, dtype=np.float32), state=np.empty(B, dtype=np.uint8),
                   idx=np.empty(B, dtype=np.int32), flag=np.empty(B, dtype=np.uint8), target=np.empty((B, 2)),
                   iters=np.empty(B, dtype=np.uint32))
        _check(lib().bmpc_fleet_tick_tobi(self._h, B, C.byref(cfg), _p(velocity, _fp), _p(vm), _p(dist, _fp), P, _p(npts, _i32p),
                                          *[_p(a, _fp) for a in arrs], si, sp, _p(out["steering_front"]), _p(out["steering_rear"]),
                                          _p(out["target_speed"], _fp), _p(out["target_distance"], _fp), _p(out["state"], _u8p),
                                          _p(out["idx"], _i32p), _p(out["flag"], _u8p), _p(out["target"]), _p(out["iters"], _u32p)))
        return out

    @staticmethod
    def _poly_strides(a, layout, B):
        if layout == "shared":
            return a.shape[0], 0, 1
        if layout == "pb":
            return a.shape[0], 1, B
        if layout == "bp":
            return a.shape[1], a.shape[1], 1
        raise ValueError(layout)

    def fleet_tick_pid(self, velocity, dist, px, py, dx, dy, vel, layout: str = "pb", npoints=None):
        """The non-MPC branch of cycle() (src/trajectory_point_follower.cpp:214-225) for len(velocity) vehicles."""
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, dx, dy, vel)]
        velocity = np.ascontiguousarray(velocity, dtype=np.float32)
        dist = None if dist is None else np.ascontiguousarray(dist, dtype=np.float32)
        B = velocity.shape[0]
        P, si, sp = self._poly_strides(arrs[0], layout, B)
        npts = None if npoints is None else np.ascontiguousarray(npoints, dtype=np.int32)
        out = dict(steering_front=np.empty(B), steering_rear=np.empty(B), target_speed=np.empty(B, dtype=np.float32),
                   target_distance=np.empty(B, dtype=np.float32), idx=np.empty(B, dtype=np.int32), flag=np.empty(B, dtype=np.uint8))
        _check(lib().bmpc_fleet_tick_pid(self._h, B, _p(velocity, _fp), _p(dist, _fp), P, _p(npts, _i32p),
                                         *[_p(a, _fp) for a in arrs], si, sp, _p(out["steering_front"]), _p(out["steering_rear"]),
                                         _p(out["target_speed"], _fp), _p(out["target_distance"], _fp),
                                         _p(out["idx"], _i32p), _p(out["flag"], _u8p)))
        return out

    def set_lookup(self, which: int, vx, vy):
        """lms::math::LookupTable<float, ASC> samples (LOOKUP_DISTANCE: src/...:66; LOOKUP_VELOCITY: :323); empty = none."""
        vx = np.ascontiguousarray(vx, dtype=np.float32); vy = np.ascontiguousarray(vy, dtype=np.float32)
        _check(lib().bmpc_set_lookup(self._h, which, len(vx), _p(vx, _fp), _p(vy, _fp)))

    def lookup(self, which: int, x):
        x = np.ascontiguousarray(x, dtype=np.float32); y = np.empty_like(x)
        _check(lib().bmpc_lookup(self._h, which, x.shape[0], _p(x, _fp), _p(y, _fp)))
        return y

    def set_drive_mode(self, foh: bool, regelpunkt_min: float = 0.6, regelpunkt_slope: float = 0.1):
        _check(lib().bmpc_fleet_set_drive_mode(self._h, int(foh), regelpunkt_min, regelpunkt_slope))

    def set_crossing(self, cfg):
        """CrossingConfig, or None to switch the stop-at-crossing clamp off."""
        _check(lib().bmpc_fleet_set_crossing(self._h, None if cfg is None else C.byref(cfg)))

    def set_pid(self, which: int, kp, ki, kd, dt):
        _check(lib().bmpc_fleet_set_pid(self._h, which, kp, ki, kd, dt))

    def reset_pid(self, which: int):
        _check(lib().bmpc_fleet_reset_pid(self._h, which))

    def get_pid(self, which: int, count: int):
        i = np.empty(count, dtype=np.float32); p = np.empty(count, dtype=np.float32); f = np.empty(count, dtype=np.uint8)
        _check(lib().bmpc_fleet_get_pid(self._h, which, count, _p(i, _fp), _p(p, _fp), _p(f, _u8p)))
        return i, p, f

    def put_pid(self, which: int, integral, prev, first):
        i = np.ascontiguousarray(integral, dtype=np.float32); p = np.ascontiguousarray(prev, dtype=np.float32)
        f = np.ascontiguousarray(first, dtype=np.uint8)
        _check(lib().bmpc_fleet_put_pid(self._h, which, i.shape[0], _p(i, _fp), _p(p, _fp), _p(f, _u8p)))

    # ---- trajectory point ---------------------------------------------------------------------------------------------------
    def extract_horizon(self, px, py, dx, dy, vel, d0, step, steps: int, npoints=None, layout: str = "shared"):
        """getTrajectoryPoint at d_t = d0 + t*step, t < steps, per instance: (idx[B][steps], flag, (px, py, dx, dy, vel))."""
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, dx, dy, vel)]
        d0 = np.ascontiguousarray(d0, dtype=np.float32); step = np.ascontiguousarray(step, dtype=np.float32)
        B = d0.shape[0]
        P, si, sp = self._poly_strides(arrs[0], layout, B)
        npts = None if npoints is None else np.ascontiguousarray(npoints, dtype=np.int32)
        idx = np.empty((B, steps), dtype=np.int32); flag = np.empty((B, steps), dtype=np.uint8)
        outs = [np.empty((B, steps), dtype=np.float32) for _ in range(5)]
        _check(lib().bmpc_extract_horizon(self._h, B, P, _p(npts, _i32p), *[_p(a, _fp) for a in arrs], si, sp, _p(d0, _fp),
                                          _p(step, _fp), steps, _p(idx, _i32p), _p(flag, _u8p), *[_p(o, _fp) for o in outs]))
        return idx, flag, tuple(outs)

    def extract_points(self, px, py, dx, dy, vel, dist, npoints=None, layout: str = "shared"):
        """getTrajectoryPoint for a batch.  layout 'shared': arrays [P] (one polyline for all); 'pb': [P][B]
        point-major (coalesced on the device); 'bp': [B][P].  Returns (idx, flag, (px, py, dx, dy, vel))."""
        arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, dx, dy, vel)]
        dist = np.ascontiguousarray(dist, dtype=np.float32)
        B = dist.shape[0]
        if layout == "shared":
            P, si, sp = arrs[0].shape[0], 0, 1
        elif layout == "pb":
            P, si, sp = arrs[0].shape[0], 1, B
        elif layout == "bp":
            P, si, sp = arrs[0].shape[1], arrs[0].shape[1], 1
        else:
            raise ValueError(layout)
        npts = None if npoints is None else np.ascontiguousarray(npoints, dtype=np.int32)
        idx = np.empty(B, dtype=np.int32); flag = np.empty(B, dtype=np.uint8)
        outs = [np.empty(B, dtype=np.float32) for _ in range(5)]
        _check(lib().bmpc_extract_points(self._h, B, P, _p(npts, _i32p), *[_p(a, _fp) for a in arrs], si, sp,
                                         _p(dist, _fp), _p(idx, _i32p), _p(flag, _u8p), *[_p(o, _fp) for o in outs]))
        return idx, flag, outs


STATIC_EXTRACT_BYTES = 4 << 20          # BMPC_STATIC_EXTRACT_BYTES (include/bmpc.h)


def extract_points_static(px, py, dx, dy, vel, dist, npoints=None, layout: str = "shared", device: int = 0, stream: int = 0):
    """BatchedMPC.extract_points without a handle and without any allocation on the device (bmpc_extract_points_static:
    the staging is a fixed array inside libbmpc.so, STATIC_EXTRACT_BYTES; larger problems raise).  Same arguments and
    results; `stream` is a cudaStream_t as an integer (0: default stream)."""
    arrs = [np.ascontiguousarray(a, dtype=np.float32) for a in (px, py, dx, dy, vel)]
    dist = np.ascontiguousarray(dist, dtype=np.float32)
    B = dist.shape[0]
    if layout == "shared":
        P, si, sp = arrs[0].shape[0], 0, 1
    elif layout == "pb":
        P, si, sp = arrs[0].shape[0], 1, B
    elif layout == "bp":
        P, si, sp = arrs[0].shape[1], arrs[0].shape[1], 1
    else:
        raise ValueError(layout)
    npts = None if npoints is None else np.ascontiguousarray(npoints, dtype=np.int32)
    idx = np.empty(B, dtype=np.int32); flag = np.empty(B, dtype=np.uint8)
    outs = [np.empty(B, dtype=np.float32) for _ in range(5)]
    _check(lib().bmpc_extract_points_static(int(device), C.c_void_p(stream or None), B, P, _p(npts, _i32p),
                                            *[_p(a, _fp) for a in arrs], si, sp, _p(dist, _fp), _p(idx, _i32p),
                                            _p(flag, _u8p), *[_p(o, _fp) for o in outs]))
    return idx, flag, outs


class ShardedMPC:
    """`capacity` controllers sharded by instance over several GPUs of this process (bmpc_group_*, include/bmpc.h):
    one handle + one host thread per device inside libbmpc, results gathered into the caller's arrays.  `devices`
    may repeat a device (two members share it)."""

    def __init__(self, S: int, I: int, N: int, capacity: int, precision: int = FP64, mode: int = EXACT, devices=None,
                 assignment: int = GROUP_INTERLEAVED):
        self.S, self.I, self.N, self.capacity = S, I, N, int(capacity)
        self._g = C.c_void_p()
        if devices is None:
            dv, nd = None, 0
        else:
            nd = len(devices); dv = (C.c_int * nd)(*devices)
        _check(lib().bmpc_group_create(C.byref(self._g), S, I, N, self.capacity, precision, mode, dv, nd, assignment))

    def close(self):
        if getattr(self, "_g", None) is not None and self._g:
            lib().bmpc_group_destroy(self._g)
            self._g = C.c_void_p()

    __del__ = close

    def __enter__(self):
        return self

    def __exit__(self, *a):
        self.close()

    def size(self) -> int:
        return lib().bmpc_group_size(self._g)

    def shard(self, k: int, count: int | None = None):
        f, n = C.c_int64(), C.c_int64()
        _check(lib().bmpc_group_shard(self._g, k, self.capacity if count is None else count, C.byref(f), C.byref(n)))
        return f.value, n.value

    def set_model(self, model, per_instance: bool = False):
        g = (lambda k: model[k]) if isinstance(model, dict) else (lambda k: getattr(model, k))
        arrs = [_c64(g(k)) for k in ("A", "B", "C", "Q", "R", "lower", "upper")]
        count = arrs[0].shape[0] if per_instance else 1
        _check(lib().bmpc_group_set_model(self._g, *[_p(a) for a in arrs], int(per_instance), count))
        return self

    def set_options(self, eps: float, max_iter: int):
        _check(lib().bmpc_group_set_options(self._g, eps, max_iter))

    def set_target(self, val, time: int = -1):
        v = _c64(val)
        per = v.ndim == 2
        _check(lib().bmpc_group_set_target(self._g, _p(v), time, int(per), v.shape[0] if per else 0))

    def reset(self):
        _check(lib().bmpc_group_reset(self._g))

    def solve(self, x0, warm: bool = False, keep_controls: bool = False, out=None):
        x = _c64(x0)
        B = x.shape[0]
        if out is None:
            u0 = np.empty((B, self.I)); it = np.empty(B, dtype=np.uint32); st = np.empty(B, dtype=np.uint8)
        else:
            u0, it, st = out
        flags = (WARM if warm else 0) | (KEEP_CONTROLS if keep_controls else 0)
        _check(lib().bmpc_group_solve(self._g, B, _p(x), _p(u0), _p(it, _u32p), _p(st, _u8p), flags))
        return u0, it, st

    __call__ = solve

    def get_controls(self, count: int | None = None, first: int = 0):
        count = self.capacity - first if count is None else count
        out = np.zeros((count, self.N, self.I))
        _check(lib().bmpc_group_get_controls(self._g, first, count, _p(out)))
        return out

    def last_kernel_ms(self) -> float:
        ms = C.c_float()
        _check(lib().bmpc_group_last_kernel_ms(self._g, C.byref(ms)))
        return ms.value
