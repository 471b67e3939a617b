#!/bin/bash
# scratch: per-call host latency of the facade's sequence + module cycle time
mkdir -p gpurun_out
for i in 1 2; do timeout 120 scripts/probes/call_latency; done | tee gpurun_out/call_latency.jsonl
timeout 120 tests/cpp/build/module_driver trajectory_controller_b200/cpp/build/libtrajectory_point_controller.so --time 2000 | tail -1 | tee gpurun_out/module_time.json
