#!/bin/bash
# 1M-instance throughput per arithmetic mode for each build tag
for tag in "$@"; do
lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so; [ "$tag" = base ] && lib=$PWD/trajectory_controller_b200/libbmpc.so
for mode in "--mode exact" "--mode fast" "--precision 32 --mode fast"; do
BMPC_LIB=$lib timeout 300 python bench.py --steps 3 --warmup 2 --no-cpu --no-fast --batch 1048576 $mode > gpurun_out/big_$tag.json 2> gpurun_out/big_$tag.err
python -c "
import json; d=json.load(open('gpurun_out/big_$tag.json')); print('$tag','$mode','value %.4e'%d['value'],'ms %.2f'%d['ms_per_step'])"
done; done
