./tests/cpp/build/test_group 2>&1 | tail -3
timeout 600 python -m pytest tests/test_group.py -m gpu -x -q 2>&1 | tail -2
timeout 600 python scripts/bench_group.py > gpurun_out/r2_group.json 2> gpurun_out/group.err || tail -5 gpurun_out/group.err; cat gpurun_out/r2_group.json | cut -c1-900
