for tag in "$@"; do
lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so
for b in 65536 131072; do
BMPC_LIB=$lib timeout 300 python bench.py --steps 5 --warmup 3 --no-cpu --batch $b > gpurun_out/pp_${tag}_$b.json 2> gpurun_out/pp_${tag}_$b.err
python -c "
import json; d=json.load(open('gpurun_out/pp_${tag}_$b.json')); print('$tag',$b,'value %.4e'%d['value'],'ms %.3f'%d['ms_per_step'])"
done; done
