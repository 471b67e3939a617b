#!/bin/bash
# usage: bash scripts/r2_scale.sh N  — bench.py on N GPUs of this box (the driver's SCALE launch), both arms
N=$1
mkdir -p gpurun_out
run() { if [ "$N" = 1 ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29519 "$@"; fi; }
timeout 600 bash -c "$(declare -f run); N=$N; run bench.py --gpus $N --steps 20 --warmup 3" > gpurun_out/r2_scale_v9_n$N.json 2> gpurun_out/scale_n$N.err; echo "bench rc=$?"
cut -c1-200 gpurun_out/r2_scale_v9_n$N.json
python - $N <<'PY'
import json,sys
d=json.load(open('gpurun_out/r2_scale_v9_n%s.json'%sys.argv[1]))
print('N',d['n_gpus'],'value %.4e e2e %.4e'%(d['value'],d['e2e']['value']),'gather',d['e2e'].get('gather'),'clocks',d['clocks'])
PY
