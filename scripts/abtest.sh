#!/bin/bash
# For each tag: parity subset (shape 4,1,30 only — QUICK builds hold nothing else) then the A/B bench.
for tag in "$@"; do
  lib=$PWD/trajectory_controller_b200/libbmpc_$tag.so
  BMPC_LIB=$lib timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -x -q -k "bike_cold_matches_dlib and 30 or closed_loop_with_moving or config2_full or 4-1-30" 2>&1 | tail -3
done
bash scripts/ab.sh "$@"
