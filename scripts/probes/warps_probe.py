"""Developer probe: kernel time of a cold (4,1,30) batch (config 2 / config 3 inputs) under the current env knobs.
usage: python scripts/probes/warps_probe.py COUNT [COUNT ...]"""
import os, sys
sys.path.insert(0, '.')
import numpy as np
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
out = []
for B in map(int, sys.argv[1:]):
    x0 = W.random_offsets_x0(B, 42)
    with BatchedMPC(4, 1, 30, B) as e:
        e.set_model(W.bicycle4_model())
        e.upload_x0(x0)
        ms = []
        for _ in range(5):
            e.solve_resident_async(B); e.sync(); ms.append(e.last_kernel_ms())
        a, b = e.last_stage_ms()
        out.append("%d: %.3f ms (A %.3f, B+C %.3f) %.2f M/s" % (B, np.median(ms[1:]), a, b, B / np.median(ms[1:]) / 1e3))
print("PG_WARPS=%s THR=%s | " % (os.environ.get("BMPC_PG_WARPS", "-"), os.environ.get("BMPC_PAIR_THRESHOLD", "-")) + " | ".join(out))
