timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -6 gpurun_out/pytest_gpu.log
for cm in 1000000000 32768; do
BMPC_TICK_CHUNK_MIN=$cm timeout 600 python scripts/bench_configs.py tick > gpurun_out/tick_$cm.json 2> gpurun_out/tick.err || tail -5 gpurun_out/tick.err
python - $cm <<'PY'
import json,sys
d=json.load(open('gpurun_out/tick_%s.json'%sys.argv[1]))
print('chunk_min',sys.argv[1],'ms_per_fleet_tick %.3f'%d['ms_per_fleet_tick'],'vehicle_ticks/s %.3e'%d['vehicle_ticks_per_s'],'pinned',d['caller_buffers_pinned'],'launches',d['kernel_launches_per_tick'])
PY
done
