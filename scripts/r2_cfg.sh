#!/bin/bash
# usage: bash scripts/r2_cfg.sh N — configs 3/4/5 (scripts/bench_multi.py all) on N GPUs of this box
N=$1
mkdir -p gpurun_out
run() { if [ "$N" = 1 ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 "$@"; fi; }
timeout 1200 bash -c "$(declare -f run); N=$N; run scripts/bench_multi.py all --out gpurun_out/r2" > gpurun_out/multi_n$N.jsonl 2> gpurun_out/multi_n$N.err; echo "bench_multi rc=$?"
tail -2 gpurun_out/multi_n$N.err | cut -c1-200
python - $N <<'PY'
import json,sys
n=sys.argv[1]
t=json.load(open(f'gpurun_out/r2_cfg3_n{n}.json'))['results']; s=json.load(open(f'gpurun_out/r2_cfg4_n{n}.json'))['results']; c=json.load(open(f'gpurun_out/r2_cfg5_n{n}.json'))
print('N=%s cfg3 exact64 %.4e fast64 %.4e ex32 %.4e f32 %.4e | cfg4 N10 %.3e N30 %.3e N60 %.3e N100 %.3e | cfg5 %.4e solves/s %.2f ms/tick gather %.1f ms'%(n,t['fp64_exact']['solves_per_s'],t['fp64_fast']['solves_per_s'],t['fp32_exact']['solves_per_s'],t['fp32_fast']['solves_per_s'],s['N10']['solves_per_s'],s['N30']['solves_per_s'],s['N60']['solves_per_s'],s['N100']['solves_per_s'],c['solves_per_s'],c['ms_per_tick'],c['gather_ms']))
PY
