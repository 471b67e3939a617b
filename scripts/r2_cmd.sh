for th in 2368 4736 9472 14208 18944; do
BMPC_PAIR_THRESHOLD=$th BMPC_PAIR_DIRECT=9472 timeout 300 python - <<'PY'
import numpy as np, os
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
x0=W.random_offsets_x0(65536,42)
out=[]
with BatchedMPC(4,1,30,65536) as e:
    e.set_model(W.bicycle4_model())
    for B in (8192,12000,16384,32768,65536):
        e.upload_x0(x0[:B]); ms=[]
        for _ in range(5):
            e.solve_resident_async(B); e.sync(); ms.append(e.last_kernel_ms())
        out.append('%d: %.3f ms'%(B,min(ms[1:])))
print('threshold',os.environ['BMPC_PAIR_THRESHOLD'],' | '.join(out))
PY
done
