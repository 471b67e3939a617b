#!/bin/bash
# usage: bash scripts/r2_multi.sh N   — configs 3/4/5 (scripts/bench_multi.py all) and bench.py on N GPUs of this box; group tests
N=$1
mkdir -p gpurun_out
run() { if [ "$N" = 1 ]; then python "$@"; else python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 "$@"; fi; }
nvidia-smi --query-gpu=index,name,clocks.max.sm --format=csv > gpurun_out/smi_n$N.txt 2>&1
./tests/cpp/build/test_group 2>&1 | tail -5 | tee gpurun_out/test_group_n$N.log
timeout 600 python -m pytest tests/test_group.py -m gpu -x -q 2>&1 | tail -3 | tee -a gpurun_out/test_group_n$N.log
timeout 1200 bash -c "$(declare -f run); N=$N; run scripts/bench_multi.py all --out gpurun_out/r2" > gpurun_out/multi_n$N.jsonl 2> gpurun_out/multi_n$N.err; echo "bench_multi rc=$?"
tail -3 gpurun_out/multi_n$N.err
timeout 600 bash -c "$(declare -f run); N=$N; run bench.py --gpus $N --steps 10 --warmup 3" > gpurun_out/r2_scale_v7_n$N.json 2> gpurun_out/scale_n$N.err; echo "bench rc=$?"
cut -c1-300 gpurun_out/r2_scale_v7_n$N.json
ls gpurun_out
