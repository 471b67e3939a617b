#!/bin/bash
# Round-2 first GPU session: full parity suite, bench line, split-setting A/B at 65,536 and 1M, small batches.
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_exact64.json 2> gpurun_out/bench_exact64.err; echo "bench rc=$?"
bash scripts/r2_ab.sh "" "base:1 base:2 base:3" 2>&1 | tee gpurun_out/ab.log
