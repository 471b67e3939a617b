timeout 600 python scripts/stress_stages.py > gpurun_out/stress.log 2>&1; echo "stress rc=$?"; grep -c " ok " gpurun_out/stress.log; grep -v " ok " gpurun_out/stress.log | tail -5
