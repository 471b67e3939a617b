#!/bin/bash
# Round-2 A/B: parity subset + bench + small batches for QUICK builds libbmpc_<tag>.so ("base" = libbmpc.so).
# usage: bash scripts/r2_ab.sh "<tags for parity>" "<tag[:split] ...  for the bench>"
mkdir -p gpurun_out
for tag in $1; do
  lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so
  echo "== parity $tag"
  BMPC_LIB=$lib timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bike_cold_matches_dlib and 30 or closed_loop_with_moving or config2_full or 4-1-30 or split or slicing" 2>&1 | tail -4
done
for spec in $2; do
  tag=${spec%%:*}; split=""; [ "$spec" != "$tag" ] && split=${spec##*:}
  lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so; [ "$tag" = base ] && lib=$PWD/trajectory_controller_b200/libbmpc.so
  for b in 65536 1048576; do
    BMPC_SPLIT=$split BMPC_LIB=$lib timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu --no-fast --batch $b > gpurun_out/ab_${tag}_${split}_$b.json 2> gpurun_out/ab_${tag}_${split}_$b.err
    python - "$spec" $b gpurun_out/ab_${tag}_${split}_$b <<'PY'
import json,sys
t,b,f=sys.argv[1:4]
try:
    d=json.load(open(f+'.json')); r=d['roofline']
    print(t,b,'value %.4e'%d['value'],'ms %.3f'%d['ms_per_step'],'stageB', r.get('kernel_ms'), 'clk',d['clocks'].get('sm_mhz'))
except Exception as e:
    print(t,b,'FAILED',e,open(f+'.err').read()[-600:])
PY
  done
  BMPC_SPLIT=$split BMPC_LIB=$lib timeout 300 python - <<'PY'
import numpy as np, os
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
x0=W.random_offsets_x0(65536,42)
out=[]
with BatchedMPC(4,1,30,65536) as e:
    e.set_model(W.bicycle4_model())
    for B in (1,32,256,4096,16384,32768):
        e.upload_x0(x0[:B]); ms=[]
        for _ in range(4):
            e.solve_resident_async(B); e.sync(); ms.append(e.last_kernel_ms())
        out.append('%d: %.3f ms'%(B,min(ms[1:])))
print('  small', ' | '.join(out))
PY
done
