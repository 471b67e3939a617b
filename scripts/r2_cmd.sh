#!/bin/bash
# scratch: the no-allocation entry point alone in a fresh process, then the whole round session
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_zz_no_alloc.py -x -q > gpurun_out/pytest_no_alloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/pytest_no_alloc_alone.log
tail -3 gpurun_out/pytest_no_alloc_alone.log
bash scripts/gpu_round.sh
tail -3 gpurun_out/pytest_gpu.log; cut -c1-400 gpurun_out/bench_exact64.json
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?"; tail -3 gpurun_out/smoke.log
