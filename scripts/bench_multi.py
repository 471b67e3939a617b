#!/usr/bin/env python
"""BASELINE.json configs[2], [3], [4] at 1/2/4/8 GPUs: a FIXED global problem is
sharded by instance over the ranks (strong scaling), one process per GPU, and the job's only collective — the final
result gather (ONE NCCL all_gather of every rank's result block u0 | iteration counts | status out of HBM, then a D2H of
the gathered job into pinned host memory on rank 0) — is INSIDE the timed region.  One JSON line per run (rank 0).

    python scripts/bench_multi.py tolerance|sweep|closed                                     # one GPU
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        scripts/bench_multi.py tolerance|sweep|closed [--batch ...]                          # N GPUs

  tolerance  config 3: 1,048,576 cold (4,1,30) instances (seed 43) in FP64 EXACT / FP64 FAST / FP32 EXACT / FP32 FAST;
             deltas of u0 and of the iteration counts against FP64 EXACT on the WHOLE gathered batch
  sweep      config 4: N in {10,30,60,100} x 64 weight combos x 1,024 x0 (seed 44), per-instance models, FP64 EXACT;
             round-robin instance assignment (the combos are blocked and differ 100x in iteration counts)
  closed     config 5: 262,144 vehicles x 500 ticks on the device, FP64 EXACT; gather of the final poses

Timing: the region runs from pinned HOST inputs (H2D of the rank's shard inside) to the gathered results in HBM on every
rank and in pinned host memory on rank 0; barrier + cudaDeviceSynchronize on both sides, wall clock, MAX over ranks; the kernels'
share is reported from CUDA events.  Clocks and throttle reasons are sampled during the timed region."""
from __future__ import annotations

import argparse
import json
import os
import sys
import time

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from bench import ClockSampler  # noqa: E402
from trajectory_controller_b200 import workloads as W  # noqa: E402
from trajectory_controller_b200.engine import EXACT, FAST, FP32, FP64, BatchedMPC  # noqa: E402


class Job:
    """Process-group plumbing of one rank."""

    def __init__(self):
        import torch
        import torch.distributed as dist
        self.torch, self.dist = torch, dist
        self.rank = int(os.environ.get("RANK", "0")); self.world = int(os.environ.get("WORLD_SIZE", "1"))
        self.local = int(os.environ.get("LOCAL_RANK", "0"))
        torch.cuda.set_device(self.local)
        self.dev = torch.device("cuda", self.local)
        if self.world > 1:
            os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
            os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")
            dist.init_process_group("nccl", device_id=self.dev)

    def barrier(self):
        if self.world > 1:
            self.dist.barrier()
        self.torch.cuda.synchronize()

    def reduce(self, x: float, op="max") -> float:
        if self.world == 1:
            return x
        t = self.torch.tensor([x], dtype=self.torch.float64, device=self.dev)
        self.dist.all_reduce(t, op=self.dist.ReduceOp.MAX if op == "max" else self.dist.ReduceOp.SUM)
        return float(t.item())

    def pinned(self, a):
        t = self.torch.empty(a.shape, dtype=self.torch.from_numpy(a[:0]).dtype).pin_memory()
        t.numpy()[...] = a
        return t

    def sampler(self):
        """Clock / throttle sampler of the job: rank 0 only (one sampler per job is enough; NVML in-process, bench.ClockSampler)."""
        if self.rank != 0:
            return None
        sm = ClockSampler(self.local, period=0.05)
        sm.start()
        return sm

    def close(self):
        if self.world > 1:
            self.dist.destroy_process_group()


class Gather:
    """The job's only collective: ONE all_gather of every rank's result block (u0 [I][cap] | iters | status live in one
    device allocation, bmpc_device_results_block) — afterwards every rank holds the whole job's results in HBM — plus one
    D2H of the gathered blocks into pinned host memory on rank 0.  Rank r holds the global instances r, r + world, ...
    (round robin): row-major [cap][world] of the gathered [world][cap] arrays is global instance order."""

    def __init__(self, job: Job, eng: BatchedMPC, count: int):
        t = job.torch
        self.job, self.count, self.world = job, count, job.world
        view, self.off_it, self.off_st = eng.device_results_block()
        self.dblk = t.as_tensor(view, device=job.dev)
        n = self.dblk.numel()
        self.g = t.empty((job.world, n), dtype=t.uint8, device=job.dev)
        self.h = t.empty((job.world, n), dtype=t.uint8).pin_memory()
        self.bytes_out = job.world * n
        self.cap, self.I = eng.capacity, eng.I
        self.ftype = np.float64 if eng.precision == FP64 else np.float32

    def __call__(self):
        if self.world > 1:
            self.job.dist.all_gather_into_tensor(self.g.view(-1), self.dblk)
        else:
            self.g[0].copy_(self.dblk)
        if self.job.rank == 0:
            self.h.copy_(self.g, non_blocking=True)

    def results(self):
        """(u0[count][I] float64, iters[count], status[count]) in global instance order (outside the timed region)."""
        self.job.torch.cuda.synchronize()
        blk = self.h.numpy() if self.job.rank == 0 else self.g.cpu().numpy()
        cap, I, w = self.cap, self.I, self.world
        es = np.dtype(self.ftype).itemsize
        u0 = blk[:, : I * cap * es].copy().view(self.ftype).reshape(w, I, cap).astype(np.float64)
        u0 = np.ascontiguousarray(u0.transpose(2, 0, 1)).reshape(cap * w, I)[: self.count]
        it = blk[:, self.off_it: self.off_it + 4 * cap].copy().view(np.uint32).reshape(w, cap)
        it = np.ascontiguousarray(it.T).reshape(-1)[: self.count].astype(np.int64)
        st = np.ascontiguousarray(blk[:, self.off_st: self.off_st + cap].T).reshape(-1)[: self.count]
        return u0, it, st


def timed_solves(job: Job, eng: BatchedMPC, x_mine, gather: Gather, reps: int):
    """reps x (bmpc_upload_x0 from pinned host + solve + gather); returns (seconds per rep = max over ranks, kernel ms)."""
    n = x_mine.shape[0]
    eng.upload_x0(x_mine); eng.solve_resident_async(n); gather(); job.barrier()          # warm-up
    t0 = time.perf_counter()
    kms = []
    for _ in range(reps):
        eng.upload_x0(x_mine)
        eng.solve_resident_async(n)
        gather()
        job.torch.cuda.synchronize()
        kms.append(eng.last_kernel_ms())
    dt = time.perf_counter() - t0
    job.barrier()
    return job.reduce(dt) / reps, job.reduce(float(np.mean(kms)))


def tolerance(job: Job, args):
    B, world, rank = args.batch, job.world, job.rank
    x0 = W.random_offsets_x0(B, 43)
    mine = job.pinned(np.ascontiguousarray(x0[rank::world])).numpy()
    cap = (B + world - 1) // world
    bike = W.bicycle4_model()
    res, ref = {}, None
    sampler = job.sampler()
    for name, prec, mode in (("fp64_exact", FP64, EXACT), ("fp64_fast", FP64, FAST), ("fp32_exact", FP32, EXACT), ("fp32_fast", FP32, FAST)):
        with BatchedMPC(4, 1, 30, cap, prec, mode, device=job.local) as eng:
            eng.set_model(bike)
            eng.set_stream(job.torch.cuda.current_stream().cuda_stream)
            g = Gather(job, eng, B)
            secs, kms = timed_solves(job, eng, mine, g, args.reps)
            u0, it, st = g.results()
        if ref is None:
            ref = (u0.copy(), it.copy())
        du = np.abs(u0 - ref[0])
        res[name] = {"solves_per_s": B / secs, "ms_per_step": secs * 1e3, "kernel_ms_slowest_rank": kms,
                     "max_abs_du": float(du.max()), "p999_abs_du": float(np.percentile(du, 99.9)), "mean_abs_du": float(du.mean()),
                     "iters_mean": float(it.mean()), "iters_equal_frac": float((it == ref[1]).mean()),
                     "max_abs_diters": int(np.abs(it - ref[1]).max()), "hit_max_iterations": int((st & 1).sum()),
                     "gather_bytes_per_rank": g.bytes_out}
    return {"config": f"configs[2]: tolerance study, {B} cold (4,1,30) instances (seed 43) sharded round-robin over {world} GPU(s); "
                      "deltas vs FP64 EXACT (== dlib bit for bit) on the whole gathered batch", "bars": {"fp64": 1e-6, "fp32": 1e-3},
            "results": res, "clocks": sampler.result() if sampler else None}


def sweep(job: Job, args):
    world, rank = job.world, job.rank
    out = {}
    sampler = job.sampler()
    for N in (10, 30, 60, 100):
        models, x0 = W.sweep_models(N, args.per_combo, 44)
        B = len(x0)
        mine_models = {k: np.ascontiguousarray(v[rank::world]) for k, v in models.items()}
        mine = job.pinned(np.ascontiguousarray(x0[rank::world])).numpy()
        cap = (B + world - 1) // world
        with BatchedMPC(4, 1, N, cap, FP64, EXACT, device=job.local) as eng:
            t0 = time.perf_counter()
            eng.set_model(mine_models, per_instance=True)
            t_model = job.reduce(time.perf_counter() - t0)
            eng.set_stream(job.torch.cuda.current_stream().cuda_stream)
            g = Gather(job, eng, B)
            secs, kms = timed_solves(job, eng, mine, g, args.reps)
            u0, it, st = g.results()
            pat = eng.pattern_name()
        flops = float(W.flops_setup(4, 1, N) * B + W.flops_per_iteration(4, 1, N) * it.sum())
        out[f"N{N}"] = {"instances": B, "pattern": pat, "solves_per_s": B / secs, "ms_per_step": secs * 1e3, "kernel_ms_slowest_rank": kms,
                        "iters_mean": float(it.mean()), "iters_p99": float(np.percentile(it, 99)), "iters_max": int(it.max()),
                        "hit_max_iterations": int((st & 1).sum()), "tflops_algorithmic": flops / secs / 1e12,
                        "set_model_s_incl_ctor_kernel_and_h2d": t_model, "u0_checksum": float(np.abs(u0).sum())}
    return {"config": f"configs[3]: horizon x (w_y, w_psi, R) grid, 64 combos x {args.per_combo} x0 (seed 44), per-instance models "
                      f"(constructor math on the device), FP64 EXACT, sharded round-robin over {world} GPU(s)", "results": out,
            "clocks": sampler.result() if sampler else None}


def closed(job: Job, args):
    from trajectory_controller_b200 import sharding
    count, ticks, world, rank = args.vehicles, args.ticks, job.world, job.rank
    path = W.sine_world_path()
    pose0 = W.closed_loop_poses(count, 45, path)
    a, b = sharding.shard_range(count, rank, world)
    n = b - a
    sampler = job.sampler()
    with BatchedMPC(4, 1, 30, n, FP64, EXACT, device=job.local) as eng:
        eng.set_model(W.bicycle4_model())
        eng.cl_init(path, pose0[a:b])
        eng.cl_run(5); eng.sync()
        l0 = eng.launch_count()
        job.barrier()
        t0 = time.perf_counter()
        eng.cl_run(ticks - 5); eng.sync()
        t_run = time.perf_counter() - t0
        r = eng.cl_get()
        pose, itot, imax = sharding.gather_results([r["pose"], r["iters_total"].astype(np.int64), r["iters_max"].astype(np.int64)], count)
        dt = time.perf_counter() - t0
        job.barrier()
        launches = eng.launch_count() - l0
    dt, t_run = job.reduce(dt), job.reduce(t_run)
    sub = pose[::64]
    d = np.min(np.hypot(sub[:, 0, None] - path[0][None, :], sub[:, 1, None] - path[1][None, :]), axis=1)
    nsolves = count * (ticks - 5)
    return {"config": f"configs[4]: closed loop, {count} vehicles x {ticks} ticks (5 warm-up ticks untimed) sharded over {world} GPU(s): path "
                      "sampling + point extraction + target + warm solve + plant on the device, FP64 EXACT; final gather of poses and "
                      "iteration totals (NCCL) inside the timed region",
            "solves_per_s": nsolves / dt, "ms_per_tick": dt / (ticks - 5) * 1e3, "ms_per_tick_without_gather": t_run / (ticks - 5) * 1e3,
            "gather_ms": (dt - t_run) * 1e3, "kernel_launches_per_rank": int(launches),
            "iters_mean_per_solve": float(itot.sum() / (count * ticks)), "iters_max": int(imax.max()),
            "max_distance_to_path_m": float(d.max()), "pose_checksum": float(np.abs(pose).sum()), "clocks": sampler.result() if sampler else None}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("what", help="tolerance | sweep | closed | all, or a comma-separated list (one process, one NCCL init)")
    ap.add_argument("--batch", type=int, default=1048576)
    ap.add_argument("--per-combo", type=int, default=1024)
    ap.add_argument("--vehicles", type=int, default=262144)
    ap.add_argument("--ticks", type=int, default=500)
    ap.add_argument("--reps", type=int, default=3)
    ap.add_argument("--out", default="", help="also write each line to <out>_cfg{3,4,5}_n<N>.json")
    args = ap.parse_args()
    runs = {"tolerance": (tolerance, 3), "sweep": (sweep, 4), "closed": (closed, 5)}
    names = list(runs) if args.what == "all" else args.what.split(",")
    saved = os.dup(1); os.dup2(2, 1)                      # stdout carries exactly one JSON line per measurement
    job = Job()
    for name in names:
        fn, cfg = runs[name]
        line = fn(job, args)
        line["n_gpus"] = job.world
        if job.rank == 0:
            os.write(saved, (json.dumps(line) + "\n").encode())
            if args.out:
                with open(f"{args.out}_cfg{cfg}_n{job.world}.json", "w") as f:
                    f.write(json.dumps(line) + "\n")
    job.close()


if __name__ == "__main__":
    main()
