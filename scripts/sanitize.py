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
print("sanitize.py done")
