#!/bin/bash
# A/B for long horizons: parity subset at N=60 and the sweep timing for each build tag
for tag in "$@"; do
lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so; [ "$tag" = base ] && lib=$PWD/trajectory_controller_b200/libbmpc.so
BMPC_LIB=$lib timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bike_cold_matches_dlib" 2>&1 | tail -2
BMPC_LIB=$lib timeout 600 python - <<'PY'
import numpy as np, time, os
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
for N in (60,):
    x0=W.random_offsets_x0(16384,44)
    with BatchedMPC(4,1,N,len(x0)) as e:
        e.set_model(W.bicycle4_model()); e.upload_x0(x0)
        for _ in range(2):
            e.solve_resident_async(len(x0)); e.sync()
        ms=e.last_kernel_ms(); u,it,st=e.download(len(x0))
        print(os.environ['BMPC_LIB'][-14:], 'N',N,'ms %.2f'%ms,'solves/s %.0f'%(len(x0)/ms*1e3),'iters',it.mean(), e.kernel_info())
PY
done
