#!/bin/bash
# scratch: small-call shortcuts A/B (BMPC_SMALL bits) + parity of the paths that use them
mkdir -p gpurun_out; : > gpurun_out/call_latency_ab.jsonl
for m in 0 1 3 7 15 0 15; do echo -n "{\"BMPC_SMALL\": $m, \"r\": " >> gpurun_out/call_latency_ab.jsonl; BMPC_SMALL=$m timeout 120 scripts/probes/call_latency | tr -d '\n' >> gpurun_out/call_latency_ab.jsonl; echo "}" >> gpurun_out/call_latency_ab.jsonl; done
cat gpurun_out/call_latency_ab.jsonl
for so in trajectory_controller_b200/cpp/build/libtrajectory_point_controller.so tests/cpp/build/libref_module_bmpc.so; do
  timeout 120 tests/cpp/build/module_driver $so --time 2000 | tail -1 | tee -a gpurun_out/module_time_ab.json
done
timeout 600 python -m pytest tests/test_cpp_boundary.py tests/test_gpu_parity.py tests/test_zz_no_alloc.py -m gpu -x -q 2>&1 | tail -5 | tee gpurun_out/pytest_small.log
python - <<'PY'
import numpy as np, time
from trajectory_controller_b200.engine import BatchedMPC
from trajectory_controller_b200 import workloads as W
x0 = W.random_offsets_x0(256, 42)
with BatchedMPC(4, 1, 30, 1) as e:
    e.set_model(W.bicycle4_model())
    ts = []
    for i in range(300):
        t = time.perf_counter(); e.solve(x0[i % 256][None, :]); ts.append(time.perf_counter() - t)
    print("batch1 (4,1,30) p50 us", np.median(ts[50:]) * 1e6)
PY
