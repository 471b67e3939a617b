#!/bin/bash
# scratch: stage B with fewer warps per CTA
mkdir -p gpurun_out; : > gpurun_out/warps.txt
for w in 8 4 5 6 3; do BMPC_PG_WARPS=$w timeout 200 python scripts/probes/warps_probe.py 65536 262144 1048576 >> gpurun_out/warps.txt 2>&1; done
for t in 4736 9472 18944; do BMPC_PG_WARPS=4 BMPC_PAIR_THRESHOLD=$t timeout 200 python scripts/probes/warps_probe.py 65536 >> gpurun_out/warps.txt 2>&1; done
cat gpurun_out/warps.txt
