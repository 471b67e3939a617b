#!/usr/bin/env python
"""Secondary measurements for BASELINE.json configs[2..4] on ONE GPU.
These are not bench.py lines (the driver's contract line stays configs[1]); each prints one JSON line that
DESIGN.md §8 quotes.  Sizes are this rank's shard of the config (the path shards by instance).

    python scripts/bench_configs.py tolerance [--batch 1048576]      # config 3: FP32 / FAST vs FP64 EXACT
    python scripts/bench_configs.py sweep     [--per-combo 1024]     # config 4: horizon x weight grid, per-instance models
    python scripts/bench_configs.py closed    [--vehicles 262144 --ticks 500]   # config 5
    python scripts/bench_configs.py tick      [--vehicles 262144]               # the reference's own (2,2,4) cycle() tick, batched
"""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from trajectory_controller_b200 import workloads as W  # noqa: E402
from trajectory_controller_b200.engine import EXACT, FAST, FP32, FP64, BatchedMPC  # noqa: E402


def timed_resident(eng, B, reps=3):
    ms = []
    for _ in range(reps):
        eng.solve_resident_async(B); eng.sync()
        ms.append(eng.last_kernel_ms())
    return float(np.median(ms))


def tolerance(args):
    B = args.batch
    x0 = W.random_offsets_x0(B, 43)
    bike = W.bicycle4_model()
    res = {}
    ref_u = ref_it = None
    for name, prec, mode in (("fp64_exact", FP64, EXACT), ("fp64_fast", FP64, FAST), ("fp32_exact", FP32, EXACT), ("fp32_fast", FP32, FAST)):
        with BatchedMPC(4, 1, 30, B, prec, mode) as eng:
            eng.set_model(bike)
            eng.upload_x0(x0)
            timed_resident(eng, B, 1)
            ms = timed_resident(eng, B, 3)
            u0, it, st = eng.download(B)
        it = it.astype(np.int64)
        if ref_u is None:
            ref_u, ref_it = u0.copy(), it.copy()
        du = np.abs(u0 - ref_u)
        res[name] = {"solves_per_s": B / ms * 1e3, "kernel_ms": ms, "max_abs_du": float(du.max()), "mean_abs_du": float(du.mean()),
                     "p999_abs_du": float(np.percentile(du, 99.9)), "iters_mean": float(it.mean()),
                     "iters_equal_frac": float((it == ref_it).mean()), "max_abs_diters": int(np.abs(it - ref_it).max()),
                     "hit_max_iterations": int((st != 0).sum())}
    print(json.dumps({"config": "configs[2]: tolerance study, %d cold (4,1,30) instances on one GPU, seed 43; deltas vs FP64 EXACT "
                                "(which equals dlib bit for bit)" % B, "bars": {"fp64": 1e-6, "fp32": 1e-3}, "results": res}))


def sweep(args):
    out = {}
    for N in (10, 30, 60, 100):
        models, x0 = W.sweep_models(N, args.per_combo, 44)
        B = len(x0)
        with BatchedMPC(4, 1, N, B, FP64, EXACT) as eng:
            t0 = time.perf_counter()
            eng.set_model(models, per_instance=True)
            t_model = time.perf_counter() - t0
            eng.upload_x0(x0)
            timed_resident(eng, B, 1)
            ms = timed_resident(eng, B, 3)
            u0, it, st = eng.download(B)
            pat = eng.pattern_name()
        it = it.astype(np.int64)
        flops = float(W.flops_setup(4, 1, N) * B + W.flops_per_iteration(4, 1, N) * it.sum())
        out[f"N{N}"] = {"instances": B, "pattern": pat, "kernel_ms": ms, "solves_per_s": B / ms * 1e3, "iters_mean": float(it.mean()),
                        "iters_p99": float(np.percentile(it, 99)), "iters_max": int(it.max()), "hit_max_iterations": int((st != 0).sum()),
                        "tflops_algorithmic": flops / ms / 1e9, "set_model_s_incl_ctor_kernel_and_h2d": t_model}
    print(json.dumps({"config": "configs[3]: horizon x (w_y, w_psi, R) grid, 64 combos x %d x0 (seed 44), per-instance models "
                                "(constructor math on the device), FP64 EXACT, one GPU" % args.per_combo, "results": out}))


def closed(args):
    count, ticks = args.vehicles, args.ticks
    path = W.sine_world_path()
    pose0 = W.closed_loop_poses(count, 45, path)
    with BatchedMPC(4, 1, 30, count, FP64, EXACT) as eng:
        eng.set_model(W.bicycle4_model())
        eng.cl_init(path, pose0)
        eng.cl_run(5); eng.sync()                       # warm-up ticks (part of the trajectory, not timed)
        l0 = eng.launch_count()
        t0 = time.perf_counter()
        eng.cl_run(ticks - 5); eng.sync()
        dt = time.perf_counter() - t0
        launches = eng.launch_count() - l0
        r = eng.cl_get()
    sub = r["pose"][::64]
    d = np.min(np.hypot(sub[:, 0, None] - path[0][None, :], sub[:, 1, None] - path[1][None, :]), axis=1)
    n = count * (ticks - 5)
    print(json.dumps({"config": "configs[4]: closed loop, %d vehicles x %d ticks (5 warm-up ticks untimed), path sampling + point "
                                "extraction + target + warm solve + plant on the device, FP64 EXACT, one GPU" % (count, ticks),
                      "solves_per_s": n / dt, "ticks_per_s": (ticks - 5) / dt, "ms_per_tick": dt / (ticks - 5) * 1e3,
                      "kernel_launches": int(launches), "iters_mean_per_solve": float(r["iters_total"].sum() / (count * ticks)),
                      "iters_max": int(r["iters_max"].max()), "max_distance_to_path_m": float(d.max())}))


def tick(args):
    """The reference's own configuration: one tobiMPC cycle() tick (dlib::mpc<2,2,4>, fresh controller per tick) for a
    fleet, through the host API (H2D of polylines and speeds, D2H of the CarCommand fields inside the timed region)."""
    B, P = args.vehicles, 48
    rng = np.random.default_rng(9)
    s = np.arange(P, dtype=np.float64)[:, None] * 0.1
    ang = rng.uniform(-0.3, 0.3, B)[None, :] + rng.uniform(-0.8, 0.8, B)[None, :] * s
    f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
    px, py = f(np.cumsum(0.1 * np.cos(ang), 0)), f(rng.uniform(-0.15, 0.15, B)[None, :] + np.cumsum(0.1 * np.sin(ang), 0))
    dx, dy, vel = f(np.cos(ang)), f(np.sin(ang)), f(np.ones((P, B)))
    speed = rng.uniform(0.2, 2.5, B).astype(np.float32)
    dist = (np.float32(0.6) + speed * np.float32(0.1)).astype(np.float32)
    pinned = False
    try:        # page-locked caller buffers: the ABI copies straight from them (pageable memory halves the PCIe rate)
        import torch
        def pin(a):
            t = torch.empty(a.shape, dtype=torch.float32).pin_memory(); t.numpy()[...] = a; return t
        keep = [pin(a) for a in (px, py, dx, dy, vel, speed, dist)]
        px, py, dx, dy, vel, speed, dist = [t.numpy() for t in keep]
        pinned = True
    except Exception:
        pass
    with BatchedMPC(2, 2, 4, B, FP64, EXACT) as eng:
        r = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel)
        l0 = eng.launch_count(); t0 = time.perf_counter()
        reps = 5
        for _ in range(reps):
            r = eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel)
        dt = (time.perf_counter() - t0) / reps
        launches = (eng.launch_count() - l0) // reps
        kernel_ms = eng.last_kernel_ms()
    out = {"config": "reference drop-in shape: tobiMPC cycle() tick, dlib::mpc<2,2,4> fresh per tick, %d vehicles, P=%d points each, "
                     "host buffers in and out (bmpc_fleet_tick_tobi), FP64 EXACT, one GPU" % (B, P),
           "ticks_per_s_per_vehicle_batch": 1 / dt, "vehicle_ticks_per_s": B / dt, "ms_per_fleet_tick": dt * 1e3,
           "caller_buffers_pinned": pinned, "solve_kernels_ms": kernel_ms, "kernel_launches_per_tick": int(launches), "iters_mean": float(r["iters"].mean()),
           "h2d_bytes_per_tick": int(5 * P * B * 4 + 2 * B * 4), "d2h_bytes_per_tick": int(B * (8 + 8 + 4 + 4 + 1 + 4 + 1 + 16 + 4))}
    # the reference's CPU cost for the same tick: constructor + cold solve of dlib::mpc<2,2,4> (the glue is negligible)
    try:
        sys.path.insert(0, os.path.join(ROOT, "tests"))
        from _oracle import DlibRef, Oracle
        n = min(B, 200000)
        v = np.where(np.abs(speed[:n].astype(np.float64)) < 0.1, 0.1, speed[:n].astype(np.float64))
        Tv = 0.1 * v; Tvl = Tv / 0.21
        A = np.zeros((n, 2, 2)); A[:, 0, 0] = 1; A[:, 0, 1] = Tv; A[:, 1, 1] = 1
        Bm = np.zeros((n, 2, 2)); Bm[:, 0, 1] = Tv; Bm[:, 1, 0] = Tvl; Bm[:, 1, 1] = -Tvl
        am = 22 * np.pi / 180
        models = dict(A=A, B=Bm, C=np.zeros((n, 2)), Q=np.tile([20.0, 7.0], (n, 1)), R=np.tile([0.0005, 10.0], (n, 1)),
                      lower=np.full((n, 2), -am), upper=np.full((n, 2), am))
        cores = len(os.sched_getaffinity(0))
        solver = DlibRef() if DlibRef.available() and DlibRef().supported(2, 2, 4) else Oracle()
        secs = solver.batch_cold(models, 4, np.zeros((n, 2)), target=r["target"][:n], per_instance=True, nthreads=cores,
                                 want_controls=False)[3 if isinstance(solver, Oracle) else 2]
        out["cpu_baseline"] = {"vehicle_ticks_per_s": n / secs, "cores": cores, "kind": "reference" if not isinstance(solver, Oracle) else "port",
                               "sample": "%d vehicles, constructor + cold solve only" % n}
    except Exception as e:      # the CPU leg is informative only
        out["cpu_baseline"] = {"error": str(e)}
    print(json.dumps(out))


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", choices=["tolerance", "sweep", "closed", "tick"])
    ap.add_argument("--batch", type=int, default=1048576)
    ap.add_argument("--per-combo", type=int, default=1024)
    ap.add_argument("--vehicles", type=int, default=262144)
    ap.add_argument("--ticks", type=int, default=500)
    a = ap.parse_args()
    {"tolerance": tolerance, "sweep": sweep, "closed": closed, "tick": tick}[a.what](a)


if __name__ == "__main__":
    main()
