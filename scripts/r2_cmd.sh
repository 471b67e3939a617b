timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" | tee -a gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu --no-fast > gpurun_out/bench_a.json 2> gpurun_out/bench_a.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.load(open('gpurun_out/bench_a.json')); r=d['roofline']
print('value %.4e e2e %.4e ms %.3f stageA %.3f stageB %.3f lat %s'%(d['value'],d['e2e']['value'],d['ms_per_step'],r['whole_solve']['stage_a_ms'],r['whole_solve']['stage_b_ms'],d['latency_us_batch1']['p50']))
PY
timeout 600 python scripts/bench_multi.py closed --ticks 105 > gpurun_out/closed_a.json 2>/dev/null; python -c "
import json; d=json.load(open('gpurun_out/closed_a.json')); print('closed ms/tick', d['ms_per_tick'], d['solves_per_s'])"
