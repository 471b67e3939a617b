timeout 1500 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
bash scripts/r2_ab2.sh "s64:15 s96:15 v2:15 s192:15 c16:15 c64:15" 2>&1 | tee gpurun_out/ab4.log
timeout 600 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_exact64.json 2> gpurun_out/bench_exact64.err; echo "bench rc=$?"
