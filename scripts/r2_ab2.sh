#!/bin/bash
# usage: bash scripts/r2_ab2.sh "<tag:split ...>"   — bench at 65,536 / 1M + batch-1 latency + small batches per spec
mkdir -p gpurun_out
for spec in $1; do
  tag=${spec%%:*}; split=""; [ "$spec" != "$tag" ] && split=${spec##*:}
  lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so; [ "$tag" = base ] && lib=$PWD/trajectory_controller_b200/libbmpc.so
  line="$spec"
  for b in 65536 1048576; do
    BMPC_SPLIT=$split BMPC_LIB=$lib timeout 300 python bench.py --steps 6 --warmup 3 --no-cpu --no-fast --batch $b > gpurun_out/ab_${tag}_${split}_$b.json 2> gpurun_out/ab_${tag}_${split}_$b.err
    line="$line | $(python - gpurun_out/ab_${tag}_${split}_$b.json <<'PY'
import json,sys
try:
    d=json.load(open(sys.argv[1])); r=d['roofline']
    print('%d: %.3e /s stageB %.3f ms lat1 %.0f us'%(d['config']['instances_per_gpu'] if 'instances_per_gpu' in d['config'] else 0, d['value'], r.get('kernel_ms') or 0, d['latency_us_batch1']['p50']))
except Exception as e:
    print('FAILED',e)
PY
)"
  done
  echo "$line"
  BMPC_SPLIT=$split BMPC_LIB=$lib timeout 300 python - <<'PY'
import numpy as np, os
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
x0=W.random_offsets_x0(65536,42)
out=[]
with BatchedMPC(4,1,30,65536) as e:
    e.set_model(W.bicycle4_model())
    for B in (1,32,4096,16384,32768):
        e.upload_x0(x0[:B]); ms=[]
        for _ in range(4):
            e.solve_resident_async(B); e.sync(); ms.append(e.last_kernel_ms())
        out.append('%d: %.3f ms'%(B,min(ms[1:])))
print('    small', ' | '.join(out))
PY
done
