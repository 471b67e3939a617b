#!/bin/bash
# Round-2 session 2: group layer tests + bench_multi smoke on one GPU
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_group.py tests/test_cpp_boundary.py -m gpu -x -q > gpurun_out/pytest_group.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_group.log
tail -15 gpurun_out/pytest_group.log
./tests/cpp/build/test_group 2>&1 | tail -8
for w in "tolerance --batch 65536 --reps 1" "sweep --per-combo 64 --reps 1" "closed --vehicles 8192 --ticks 30"; do
  timeout 600 python scripts/bench_multi.py $w > gpurun_out/multi_smoke_$(echo $w | cut -d' ' -f1).json 2> gpurun_out/multi_smoke.err || tail -20 gpurun_out/multi_smoke.err
  cut -c1-600 gpurun_out/multi_smoke_$(echo $w | cut -d' ' -f1).json
done
