"""Developer probe: one solve of a config-4 sweep slice (per-instance models) at horizon N, for `ncu --set full` of the
stage-C kernel.  usage: python scripts/sweep_probe.py N [per_combo]"""
import sys
sys.path.insert(0, '.')
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
N = int(sys.argv[1]); per = int(sys.argv[2]) if len(sys.argv) > 2 else 256
models, x0 = W.sweep_models(N, per, 44)
with BatchedMPC(4, 1, N, len(x0)) as e:
    e.set_model(models, per_instance=True)
    e.upload_x0(x0)
    for _ in range(2):
        e.solve_resident_async(len(x0)); e.sync()
    print(N, len(x0), e.last_kernel_ms(), e.last_stage_ms())
