#!/bin/bash
# kernel time of cold (4,1,30) solves for small and medium batches, per build tag
for tag in "$@"; do
lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so; [ "$tag" = base ] && lib=$PWD/trajectory_controller_b200/libbmpc.so
BMPC_LIB=$lib timeout 300 python - <<'PY'
import numpy as np, os
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
x0=W.random_offsets_x0(65536,42)
out=[]
with BatchedMPC(4,1,30,65536) as e:
    e.set_model(W.bicycle4_model())
    for B in (1,32,256,1024,4096,16384,32768,65536):
        e.upload_x0(x0[:B]); ms=[]
        for _ in range(4):
            e.solve_resident_async(B); e.sync(); ms.append(e.last_kernel_ms())
        out.append('%d: %.3f ms'%(B,min(ms[1:])))
print(os.environ['BMPC_LIB'][-14:], ' | '.join(out))
PY
done
