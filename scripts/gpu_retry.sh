#!/bin/bash
# usage: [GPUS=N] scripts/gpu_retry.sh <timeout_s> <logfile> <command...>   — retries while the pod answers "busy" (exit 3)
t=$1; log=$2; shift 2
g=""; [ -n "$GPUS" ] && g="--gpus $GPUS"
for try in $(seq 1 40); do
  /usr/local/graft/bin/gpurun $g --timeout $t -- "$@" > $log 2>&1; rc=$?
  if [ $rc -ne 3 ]; then echo "gpurun rc=$rc (try $try)" >> $log; exit $rc; fi
  sleep 120
done
