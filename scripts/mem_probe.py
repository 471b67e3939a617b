"""Developer probe: one call each of the memory-side kernels at full width (262,144 polylines x 48 points, point-major;
closed-loop prepare / plant; per-instance model upload + constructor kernel), for `ncu --set full` captures."""
import sys
import numpy as np
sys.path.insert(0, '.')
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC

B, P = 262144, 48
rng = np.random.default_rng(9)
s = np.arange(P, dtype=np.float64)[:, None] * 0.1
ang = rng.uniform(-0.3, 0.3, B)[None, :] + rng.uniform(-0.8, 0.8, B)[None, :] * s
f = lambda a: np.ascontiguousarray(a, dtype=np.float32)
px, py = f(np.cumsum(0.1 * np.cos(ang), 0)), f(rng.uniform(-0.15, 0.15, B)[None, :] + np.cumsum(0.1 * np.sin(ang), 0))
dx, dy, vel = f(np.cos(ang)), f(np.sin(ang)), f(np.ones((P, B)))
speed = rng.uniform(0.2, 2.5, B).astype(np.float32)
dist = (np.float32(0.6) + speed * np.float32(0.1)).astype(np.float32)
with BatchedMPC(2, 2, 4, B) as eng:
    for _ in range(2):
        eng.extract_points(px, py, dx, dy, vel, dist, layout="pb")
        eng.extract_horizon(px, py, dx, dy, vel, dist, np.full(B, 0.1, np.float32), 30, layout="pb")
        eng.fleet_tick_tobi(speed, dist, px, py, dx, dy, vel)
path = W.sine_world_path()
pose0 = W.closed_loop_poses(B, 45, path)
with BatchedMPC(4, 1, 30, B) as eng:
    eng.set_model(W.bicycle4_model())
    eng.cl_init(path, pose0)
    eng.cl_run(2)
    eng.cl_set_horizon_targets(True)
    eng.cl_run(2)
models, x0 = W.sweep_models(30, 1024, 44)
with BatchedMPC(4, 1, 30, len(x0)) as eng:
    eng.set_model(models, per_instance=True)
    eng.solve(x0[:4096])
print("ok")
