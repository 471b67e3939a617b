"""Developer probe: a small cold batch (lone-warp regime of the stage-C kernel) for ncu source-level stall sampling."""
import sys
import numpy as np
sys.path.insert(0, '.')
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
x0 = W.random_offsets_x0(B, 42)
with BatchedMPC(4, 1, 30, B) as e:
    e.set_model(W.bicycle4_model())
    e.upload_x0(x0)
    for _ in range(3):
        e.solve_resident_async(B); e.sync()
    print(B, e.last_kernel_ms(), e.last_stage_ms())
