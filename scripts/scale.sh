#!/bin/bash
# 1 -> 8 GPU weak-scaling run of bench.py (one process per GPU via torchrun), results in gpurun_out/scale_N.json
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/scale_gpus.txt
for n in "$@"; do
  if [ "$n" = 1 ]; then timeout 300 python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu $EXTRA > gpurun_out/scale${TAG}_$n.json 2> gpurun_out/scale${TAG}_$n.err
  else timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $n --steps 10 --warmup 3 --no-cpu $EXTRA > gpurun_out/scale${TAG}_$n.json 2> gpurun_out/scale${TAG}_$n.err; fi
  echo "N=$n rc=$?"; python -c "
import json; d=json.load(open('gpurun_out/scale${TAG}_$n.json')); print($n, 'value %.4e'%d['value'], 'e2e %.4e'%d['e2e']['value'], 'ms %.3f'%d['ms_per_step'])" || tail -5 gpurun_out/scale${TAG}_$n.err
done
