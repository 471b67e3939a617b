timeout 600 python scripts/bench_multi.py tolerance --batch 262144 --reps 1 > gpurun_out/tol_check.json 2> gpurun_out/tol.err || tail -5 gpurun_out/tol.err
python - <<'PY'
import json
d=json.load(open('gpurun_out/tol_check.json'))['results']
print({k:(round(v['solves_per_s']),v['max_abs_du'],v['iters_equal_frac'],v['iters_mean']) for k,v in d.items()})
PY
