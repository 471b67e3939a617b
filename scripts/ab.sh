#!/bin/bash
# A/B: bench.py (resident + e2e legs) for each libbmpc_<tag>.so given on the command line ("base" = libbmpc.so).
mkdir -p gpurun_out
for tag in "$@"; do
  if [ "$tag" = base ]; then lib=trajectory_controller_b200/libbmpc.so; else lib=trajectory_controller_b200/libbmpc_$tag.so; fi
  BMPC_LIB=$PWD/$lib timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu > gpurun_out/ab_$tag.json 2> gpurun_out/ab_$tag.err
  python - "$tag" <<'PY'
import json,sys
t=sys.argv[1]
try:
    d=json.load(open(f'gpurun_out/ab_{t}.json')); r=d['roofline']
    print(t, 'value %.4e'%d['value'], 'ms %.3f'%d['ms_per_step'], 'launch', r['launch'], 'iters', d['iterations']['mean'])
except Exception as e:
    print(t, 'FAILED', e, open(f'gpurun_out/ab_{t}.err').read()[-400:])
PY
done
