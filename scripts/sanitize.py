#!/usr/bin/env python
"""Small end-to-end exercise of every kernel family for compute-sanitizer runs (memcheck / racecheck / synccheck):
cold + warm two-stage solves with time slicing, per-instance models (constructor kernel), the fleet tick, point
extraction and a short closed loop.  Sizes are tiny because the sanitizer slows kernels down 10-100x.
    compute-sanitizer --tool memcheck python scripts/sanitize.py
"""
import os
import sys

import numpy as np

sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trajectory_controller_b200 import workloads as W  # noqa: E402
from trajectory_controller_b200.engine import BatchedMPC  # noqa: E402

B = int(os.environ.get("SAN_B", "1500"))
bike = W.bicycle4_model()
x0 = W.random_offsets_x0(B, 3)
with BatchedMPC(4, 1, 30, B) as eng:
    eng.set_model(bike)
    u0, it, st = eng.solve(x0, keep_controls=True)
    u0_first, it_first = u0.copy(), it.copy()
    print("cold two-stage:", it.mean(), "iterations mean,", int((it > 50).sum()), "instances reached stage B")
    eng.reset()
    for _ in range(2):
        u0, it, st = eng.solve(x0, warm=True)
    eng.set_split(0)
    u1, it1, _ = eng.solve(x0)
    print("one-kernel:", it1.mean())
models, xs = W.sweep_models(10, 8, 44)
with BatchedMPC(4, 1, 10, len(xs)) as eng:
    eng.set_model(models, per_instance=True)
    u0, it, st = eng.solve(xs)
    print("per-instance models:", it.mean())
rng = np.random.default_rng(0)
n, P = 300, 32
s = np.arange(P)[:, None] * 0.1
ang = rng.uniform(-0.3, 0.3, n)[None, :] + rng.uniform(-0.5, 0.5, n)[None, :] * s
f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
px, py = f(np.cumsum(0.1 * np.cos(ang), 0)), f(np.cumsum(0.1 * np.sin(ang), 0))
speed = rng.uniform(0, 2, n).astype(np.float32)
with BatchedMPC(2, 2, 4, n) as eng:
    r = eng.fleet_tick_tobi(speed, 0.6 + 0.1 * speed, px, py, f(np.cos(ang)), f(np.sin(ang)), f(np.ones((P, n))))
    print("fleet tick:", r["iters"].mean(), float(np.abs(r["steering_front"]).max()))
path = W.sine_world_path()
with BatchedMPC(4, 1, 30, 256) as eng:
    eng.set_model(bike)
    eng.cl_init(path, W.closed_loop_poses(256, 45, path))
    eng.cl_run(3)
    print("closed loop:", eng.cl_get()["iters_total"].mean())
# ---- round 2: stage B -> stage C hand-over (forced by a low threshold through the developer env knobs), per-instance models
# and a longer horizon in stage C, horizon extraction, closed loop with per-step targets, chunked fleet tick, groups
os.environ["BMPC_PAIR_THRESHOLD"] = "600"; os.environ["BMPC_PAIR_DIRECT"] = "100"
with BatchedMPC(4, 1, 30, B) as eng:
    eng.set_model(bike)
    u2, it2, _ = eng.solve(x0)
    print("A -> B -> C hand-over:", it2.mean(), "equal to the first solve:", bool(np.array_equal(u2, u0_first) and np.array_equal(it2, it_first)))
del os.environ["BMPC_PAIR_THRESHOLD"], os.environ["BMPC_PAIR_DIRECT"]
models60, xs60 = W.sweep_models(60, 4, 44)
with BatchedMPC(4, 1, 60, len(xs60)) as eng:
    eng.set_model(models60, per_instance=True)
    eng.set_options(0.01, 400)
    u0, it, st = eng.solve(xs60)
    print("stage C, N=60, per-instance models:", it.mean())
with BatchedMPC(4, 1, 30, n) as eng:
    d0 = (0.6 + 0.1 * speed).astype(np.float32)
    r = eng.extract_horizon(px, py, f(np.cos(ang)), f(np.sin(ang)), f(np.ones((P, n))), d0, np.full(n, 0.1, np.float32), 30, layout="pb")
    print("extract_horizon:", int(r[0].max()))
with BatchedMPC(4, 1, 30, 256) as eng:
    eng.set_model(bike)
    eng.cl_init(path, W.closed_loop_poses(256, 45, path))
    eng.cl_set_horizon_targets(True)
    eng.cl_run(3)
    print("closed loop, per-step targets:", eng.cl_get()["iters_total"].mean())
os.environ["BMPC_TICK_CHUNK_MIN"] = "1"
with BatchedMPC(2, 2, 4, n) as eng:
    r = eng.fleet_tick_tobi(speed, 0.6 + 0.1 * speed, px, py, f(np.cos(ang)), f(np.sin(ang)), f(np.ones((P, n))))
    print("chunked fleet tick:", r["iters"].mean())
from trajectory_controller_b200.engine import ShardedMPC  # noqa: E402
with ShardedMPC(4, 1, 30, 300, devices=[0, 0]) as grp:
    grp.set_model(bike)
    ug, itg, _ = grp.solve(x0[:300])
    print("group:", itg.mean())
with BatchedMPC(4, 1, 30, 4) as eng:          # small-call paths
    eng.set_model(bike); eng.set_target([0.01, 0, 0, 0])
    print("small solve:", eng.solve(x0[:1], warm=True)[1])
print("sanitize.py done")
