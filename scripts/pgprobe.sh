#!/bin/bash
# A/B helper: bench at two batch sizes + ncu --set full of the stage-B kernel for one build tag.
tag=$1
lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so; [ "$tag" = base ] && lib=$PWD/trajectory_controller_b200/libbmpc.so
mkdir -p gpurun_out
for b in 65536 1048576; do
  BMPC_LIB=$lib timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --batch $b > gpurun_out/pp_${tag}_$b.json 2> gpurun_out/pp_${tag}_$b.err
  python - "$tag" $b <<'PY'
import json,sys
t,b=sys.argv[1:3]
try:
    d=json.load(open(f'gpurun_out/pp_{t}_{b}.json')); r=d['roofline']
    print(t,b,'value %.4e'%d['value'],'ms %.3f'%d['ms_per_step'],'frac %.3f'%r['frac'],'clk',d['clocks'].get('sm_mhz'))
except Exception as e:
    print(t,b,'FAILED',e,open(f'gpurun_out/pp_{t}_{b}.err').read()[-400:])
PY
done
BMPC_LIB=$lib timeout 900 ncu --set full --clock-control none --import-source on -k regex:${2:-pg_kernel} -s 3 -c 1 -o gpurun_out/pg_$tag -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_pg_$tag.log 2>&1
ls -la gpurun_out/pg_$tag.ncu-rep
