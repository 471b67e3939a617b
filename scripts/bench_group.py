#!/usr/bin/env python
"""One host process, all GPUs of the box (bmpc_group_*, include/bmpc.h): config 3's 1,048,576 cold (4,1,30) instances
through bmpc_group_solve from HOST arrays (H2D, solve, D2H of every member overlapping; results in the caller's arrays).
One JSON line.   python scripts/bench_group.py [--batch N] [--devices K]"""
import argparse, json, os, sys, time
import numpy as np
sys.path.insert(0, os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import EXACT, FP64, GROUP_CONTIGUOUS, GROUP_INTERLEAVED, BatchedMPC, ShardedMPC

ap = argparse.ArgumentParser(); ap.add_argument("--batch", type=int, default=1 << 20); ap.add_argument("--devices", type=int, default=0)
a = ap.parse_args()
import torch
nd = a.devices or torch.cuda.device_count()
B = a.batch
x0 = W.random_offsets_x0(B, 43)
px = torch.empty((B, 4), dtype=torch.float64).pin_memory(); px.numpy()[:] = x0
out = (torch.empty((B, 1), dtype=torch.float64).pin_memory().numpy(), torch.empty(B, dtype=torch.int32).pin_memory().numpy().view(np.uint32),
       torch.empty(B, dtype=torch.uint8).pin_memory().numpy())
res = {}
ref = None
for name, asg in (("contiguous", GROUP_CONTIGUOUS), ("interleaved", GROUP_INTERLEAVED)):
    with ShardedMPC(4, 1, 30, B, FP64, EXACT, devices=list(range(nd)), assignment=asg) as g:
        g.set_model(W.bicycle4_model())
        g.solve(px.numpy(), out=out)
        t0 = time.perf_counter(); reps = 3
        for _ in range(reps):
            g.solve(px.numpy(), out=out)
        dt = (time.perf_counter() - t0) / reps
        kms = g.last_kernel_ms()
    if ref is None:
        ref = (out[0].copy(), out[1].copy())
    res[name] = {"solves_per_s": B / dt, "ms_per_call": dt * 1e3, "kernel_ms_slowest_member": kms,
                 "equal_to_first": bool(np.array_equal(out[0], ref[0]) and np.array_equal(out[1], ref[1])), "iters_mean": float(out[1].mean())}
# one handle on device 0 solving a sample: the group's results are those of a single handle
n = min(B, 65536)
with BatchedMPC(4, 1, 30, n) as one:
    one.set_model(W.bicycle4_model())
    u1, i1, _ = one.solve(x0[:n])
res["matches_single_handle_on_first_%d" % n] = bool(np.array_equal(u1, ref[0][:n]) and np.array_equal(i1, ref[1][:n]))
print(json.dumps({"config": "configs[2] batch (%d cold (4,1,30) instances, seed 43, FP64 EXACT) through bmpc_group_solve: ONE host process, %d GPUs, "
                            "pinned host arrays in and out" % (B, nd), "n_gpus": nd, "results": res}))
