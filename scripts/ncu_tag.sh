#!/bin/bash
# ncu --set full of the solve kernel for one A/B build: scripts/ncu_tag.sh <tag>
tag=$1
lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so; [ "$tag" = base ] && lib=$PWD/trajectory_controller_b200/libbmpc.so
BMPC_LIB=$lib timeout 900 ncu --set full --clock-control none --import-source on -k regex:solve_kernel -s 3 -c 1 -o gpurun_out/solve_$tag -f python bench.py --steps 2 --warmup 3 --no-cpu > gpurun_out/ncu_$tag.log 2>&1
ls -la gpurun_out/solve_$tag.ncu-rep
