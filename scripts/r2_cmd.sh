for sp in 31 15; do
BMPC_SPLIT=$sp timeout 600 python scripts/bench_multi.py sweep --reps 2 > gpurun_out/sweep_split$sp.json 2> gpurun_out/sweep_split$sp.err || tail -5 gpurun_out/sweep_split$sp.err
python - $sp <<'PY'
import json,sys
d=json.load(open('gpurun_out/sweep_split%s.json'%sys.argv[1]))['results']
print('split',sys.argv[1],{k:(round(v['solves_per_s']),round(v['kernel_ms_slowest_rank'],2),v['u0_checksum']) for k,v in d.items()})
PY
done
