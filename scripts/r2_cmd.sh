for lib in "" roll; do
  L=$PWD/trajectory_controller_b200/libbmpc.so; [ -n "$lib" ] && L=$PWD/trajectory_controller_b200/libbmpc_$lib.so
  echo "== lib ${lib:-main}"
  BMPC_LIB=$L timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_closed_loop.py -m gpu -x -q -k "bike_cold or split_solve or sweep_slice or closed_loop_matches or tolerance or config2" 2>&1 | tail -2
  BMPC_LIB=$L timeout 600 python scripts/bench_multi.py sweep --reps 2 > gpurun_out/sweep_${lib:-main}.json 2> gpurun_out/sweep.err || tail -5 gpurun_out/sweep.err
  python - ${lib:-main} <<'PY'
import json,sys
d=json.load(open('gpurun_out/sweep_%s.json'%sys.argv[1]))['results']
print('sweep',sys.argv[1],{k:(round(v['solves_per_s']),round(v['kernel_ms_slowest_rank'],2),v['u0_checksum']) for k,v in d.items()})
PY
done
