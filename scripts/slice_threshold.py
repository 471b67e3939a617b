#!/usr/bin/env python
"""Stage-B time slicing on/off (bmpc_set_split 1 / 2) per batch size — validates the host-side threshold (bmpc_solve.cu)."""
import os, sys
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
x0 = W.random_offsets_x0(262144, 42)
with BatchedMPC(4, 1, 30, 262144) as e:
    e.set_model(W.bicycle4_model())
    for B in (8192, 32768, 65536, 98304, 131072, 151552, 196608, 262144):
        e.upload_x0(x0[:B]); row = []
        for mode in (1, 2):
            e.set_split(mode); ms = []
            for _ in range(4):
                e.solve_resident_async(B); e.sync(); ms.append(e.last_kernel_ms())
            row.append(min(ms[1:]))
        print(f"{B:7d}: sliced-if-small {row[0]:.3f} ms   never sliced {row[1]:.3f} ms   ratio {row[1] / row[0]:.3f}")
