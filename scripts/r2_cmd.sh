#!/bin/bash
# scratch: stage A / stage C time of a lone (4,1,30) instance
mkdir -p gpurun_out
BMPC_SMALL=13 python - <<'PY' | tee gpurun_out/lone_stage.txt
import numpy as np, time
from trajectory_controller_b200.engine import BatchedMPC
from trajectory_controller_b200 import workloads as W
x0 = W.random_offsets_x0(64, 42)
with BatchedMPC(4, 1, 30, 1) as e:
    e.set_model(W.bicycle4_model())
    rows = []
    for i in range(64):
        for _ in range(3):
            t = time.perf_counter(); u0, it, st = e.solve(x0[i][None, :]); dt = time.perf_counter() - t
        a, b = e.last_stage_ms()
        rows.append((int(it[0]), a * 1e3, b * 1e3, dt * 1e6))
    r = np.array(rows)
    print("iters  stageA_us  stageC_us  call_us")
    for row in r[:12]: print("%5d %9.1f %9.1f %9.1f" % tuple(row))
    print("median stage A %.1f us; stage C per iteration %.3f us; call - kernels %.1f us" % (np.median(r[:,1]), np.median(r[:,2] / np.maximum(r[:,0] - 50, 1)), np.median(r[:,3] - r[:,1] - r[:,2])))
    # max_iter = 50: stage A alone
    e.set_options(0.01, 50)
    for _ in range(3): e.solve(x0[0][None, :])
    print("max_iter=50:", e.last_stage_ms(), e.last_kernel_ms())
PY
