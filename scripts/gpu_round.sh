#!/bin/bash
# One GPU-box session: parity tests, bench (4 variants), ncu launch list, one ncu --set full capture of the solve kernel.
set -x
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,clocks.sm --format=csv > gpurun_out/smi.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_exact64.json 2> gpurun_out/bench_exact64.err; echo "bench rc=$?"
timeout 300 python bench.py --steps 10 --warmup 3 --mode fast --no-cpu > gpurun_out/bench_fast64.json 2>> gpurun_out/bench_exact64.err
timeout 300 python bench.py --steps 10 --warmup 3 --precision 32 --no-cpu > gpurun_out/bench_exact32.json 2>> gpurun_out/bench_exact64.err
timeout 300 python bench.py --steps 10 --warmup 3 --precision 32 --mode fast --no-cpu > gpurun_out/bench_fast32.json 2>> gpurun_out/bench_exact64.err
timeout 300 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.json 2>> gpurun_out/bench_exact64.err
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches.csv python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/ncu_launches.log 2>&1
timeout 900 ncu --set full --clock-control none --import-source on -k 'regex:solve_kernel|pg_kernel|pair_kernel' -s 9 -c 3 -o gpurun_out/solve_full -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_full.log 2>&1
ls -la gpurun_out
timeout 600 python scripts/sanitize.py > gpurun_out/sanitize_plain.log 2>&1; tail -12 gpurun_out/sanitize_plain.log
