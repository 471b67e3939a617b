timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
for m in oracle/_ref/libref_module_dlib.so trajectory_controller_b200/cpp/build/libtrajectory_point_controller.so tests/cpp/build/libref_module_bmpc.so; do ./tests/cpp/build/module_driver $m --time 2000; done | tee gpurun_out/module_latency.jsonl
