#!/usr/bin/env python
"""bench.py — MPC solves/sec of the batched dlib::mpc-compatible engine (BASELINE.json metric), one JSON line.

    python bench.py [--gpus N] [--steps K] [--warmup W] [--impl ours|reference] [--mode exact|fast] [--precision 64|32]
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port P \
        bench.py --gpus N --steps K --warmup W

Workload (BASELINE.json configs[1], SURVEY.md §8(d) "Config 2"): 65,536 independent cold dlib::mpc<4,1,30>
controllers PER GPU (weak scaling: rank r draws its x0 with seed 42 + r), shared synthetic bicycle model, constant
target 0, eps 0.01, max_iterations 10000, FP64.  One "step" = one operator() (mpc.h:216-240) for every instance
of the batch.

  value  : solves/s with x0 resident in HBM (kernel launches only), CUDA events on the launching stream, an L2 flush
           (256 MiB memset) between timed steps, max over ranks.
  e2e    : the same metric through the C ABI call bmpc_solve() with pinned HOST buffers (H2D of x0, layout change,
           solve, D2H of u0 / iteration counts / status inside the timed region).
  roofline: the solver is bound by the FP64 CUDA-core pipe; `achieved` = algorithmic flops
           (F_setup + iters_i * F_iter summed over the instances one launch solves) / mean launch duration, `peak` = the
           FP64 FMA peak measured on this device by bmpc_probe_peak (MEASURED_PEAKS.json has no FP64 figure).
  cpu_baseline / --impl reference: the real dlib::mpc (oracle/_ref/libdlibref.so, built from the reference's own
           header) or, if that library did not travel, the C restatement (oracle/liboracle.so) — one controller per
           host thread on all host cores, on a bounded sample of the same instances.

The product path is libbmpc.so only; oracle/ is touched by the CPU legs alone.
"""
from __future__ import annotations

import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

from trajectory_controller_b200 import workloads as W  # noqa: E402

_emit = None
METRIC = "mpc_solves_per_sec_horizon30"
UNIT = "solves/s"
S_, I_, N_ = 4, 1, 30
BATCH = 65536
SEED = 42


def parse_args():
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=10)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--mode", default="exact", choices=["exact", "fast"])
    ap.add_argument("--precision", type=int, default=64, choices=[64, 32])
    ap.add_argument("--batch", type=int, default=BATCH, help="instances per GPU (default: the BASELINE config)")
    ap.add_argument("--cpu-sample", type=int, default=0, help="instances of the CPU baseline sample (0 = auto)")
    ap.add_argument("--no-cpu", action="store_true", help="skip the cpu_baseline leg (profiling runs)")
    ap.add_argument("--no-fast", action="store_true", help="skip the FAST-mode (FMA) side measurement of an exact run")
    return ap.parse_args()


def config_dict(args, world):
    return {"workload": "configs[1]: 65,536 cold dlib::mpc<4,1,30> instances per GPU, random e_y/e_psi offsets "
                        "(mt19937_64 seed 42+rank), shared bicycle4 model, target 0, eps 0.01, max_iter 10000",
            "S": S_, "I": I_, "horizon": N_, "instances_per_gpu": args.batch, "global_instances": args.batch * world,
            "precision": f"fp{args.precision}", "mode": args.mode, "sharding": f"instances x{world}",
            "l2": "flushed between timed steps (256 MiB memset); the solver reads <5 MiB per step"}


class ClockSampler(threading.Thread):
    """SM clock + throttle reasons, sampled while the timed region runs: NVML in-process (pynvml) when it loads — a
    sample costs microseconds, nothing is forked — else one `nvidia-smi` query per sample."""
    Q = "clocks.sm,clocks.max.sm,clocks_event_reasons.active"
    NAMES = {0x1: "gpu_idle", 0x2: "applications_clocks_setting", 0x4: "sw_power_cap", 0x8: "hw_slowdown",
             0x10: "sync_boost", 0x20: "sw_thermal_slowdown", 0x40: "hw_thermal_slowdown",
             0x80: "hw_power_brake_slowdown", 0x100: "display_clock_setting"}

    def __init__(self, index: int, period: float = 0.02):
        super().__init__(daemon=True)
        self.index, self.stop_flag, self.samples, self.reasons, self.maxmhz = index, threading.Event(), [], set(), None
        self.period, self.source = period, None

    def _note(self, mhz: float, max_mhz: float, bits: int):
        self.samples.append(float(mhz)); self.maxmhz = float(max_mhz)
        for b, n in self.NAMES.items():
            if bits & b and n != "gpu_idle":
                self.reasons.add(n)

    def _run_nvml(self) -> bool:
        try:
            import pynvml as nv
            nv.nvmlInit()
            dev = nv.nvmlDeviceGetHandleByIndex(self.index)
            reasons = getattr(nv, "nvmlDeviceGetCurrentClocksEventReasons", None) or nv.nvmlDeviceGetCurrentClocksThrottleReasons
            max_mhz = nv.nvmlDeviceGetMaxClockInfo(dev, nv.NVML_CLOCK_SM)
            nv.nvmlDeviceGetClockInfo(dev, nv.NVML_CLOCK_SM); reasons(dev)          # fail here, not mid-run
        except Exception:
            return False
        self.source = "nvml"
        while not self.stop_flag.is_set():
            try:
                self._note(nv.nvmlDeviceGetClockInfo(dev, nv.NVML_CLOCK_SM), max_mhz, int(reasons(dev)))
            except Exception:
                pass
            self.stop_flag.wait(self.period)
        return True

    def run(self):
        if self._run_nvml():
            return
        self.source = "nvidia-smi"
        while not self.stop_flag.is_set():
            try:
                out = subprocess.run(["nvidia-smi", "-i", str(self.index), f"--query-gpu={self.Q}",
                                      "--format=csv,noheader,nounits"], capture_output=True, text=True, timeout=5).stdout
                f = [x.strip() for x in out.strip().split(",")]
                self._note(float(f[0]), float(f[1]), int(f[2], 16) if f[2].lower().startswith("0x") else 0)
            except Exception:
                pass
            self.stop_flag.wait(max(self.period, 0.05))

    def result(self):
        self.stop_flag.set(); self.join(timeout=6)
        med = float(np.median(self.samples)) if self.samples else None
        return {"sm_mhz": med, "sm_max_mhz": self.maxmhz, "reasons": sorted(self.reasons), "samples": len(self.samples),
                "source": self.source}


# ------------------------------------------------------------------------------------------------- CPU legs
def ncu_traffic(kernel_prefix: str, build_id: str):
    """dram__bytes_read.sum + dram__bytes_write.sum per launch of the dominant kernel, from the `ncu --set full` capture
    of THIS build running this command: profiles/ncu_traffic.json maps a build id (bmpc_version(): digest of the
    kernel sources) to {kernel prefix: {"bytes": ..., "capture": file}}.  None when the running build has no capture."""
    try:
        table = json.load(open(os.path.join(ROOT, "profiles", "ncu_traffic.json")))
        ent = table.get(build_id, {}).get(kernel_prefix)
        return (float(ent["bytes"]), ent.get("capture")) if ent else (None, None)
    except Exception:
        return (None, None)


def host_cores() -> int:
    try:
        return len(os.sched_getaffinity(0))
    except AttributeError:
        return os.cpu_count() or 1


def cpu_solver():
    """(kind, callable(model, N, x0, nthreads) -> (seconds, u0, iters or None)).  The one place bench.py executes oracle/."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import DlibRef, Oracle
    if DlibRef.available():
        ref = DlibRef()
        if ref.supported(S_, I_, N_):
            def run_ref(m, N, x0, nt):
                u0, _, secs = ref.batch_cold(m, N, x0, nthreads=nt, want_controls=False)
                return secs, u0, None                      # dlib::mpc keeps its iteration counter private (mpc.h:270)
            return "reference", run_ref
    orc = Oracle()

    def run_port(m, N, x0, nt):
        u0, _, it, secs = orc.batch_cold(m, N, x0, nthreads=nt, want_controls=False)
        return secs, u0, it
    return "port", run_port


def oracle_iterations(x0, cores):
    """Iteration counts of the C restatement (pinned bit-for-bit to the real header, tests/test_oracle.py) — the
    checker for the GPU's iteration counts, which dlib::mpc itself does not expose."""
    sys.path.insert(0, os.path.join(ROOT, "tests"))
    from _oracle import Oracle
    u0, _, it, _ = Oracle().batch_cold(W.bicycle4_model(), N_, x0, nthreads=cores, want_controls=False)
    return u0, it


def cpu_baseline(sample: int, seed: int, gpu=None):
    """The timed CPU leg; with `gpu` = (u0, iters) of the same batch from the engine (EXACT FP64 runs), the solutions
    the CPU leg just computed become the in-run parity check of the bench line."""
    kind, run = cpu_solver()
    cores = host_cores()
    x0 = W.random_offsets_x0(sample, seed)
    secs, u0_cpu, it_cpu = run(W.bicycle4_model(), N_, x0, cores)
    parity = None
    if gpu is not None:
        gu0, git = gpu[0][:sample], gpu[1][:sample]
        if it_cpu is None:
            _, it_cpu = oracle_iterations(x0, cores)
        du = np.abs(gu0 - u0_cpu)
        parity = {"checked": int(sample), "max_abs_du0": float(du.max()), "u0_equal": bool(np.all(gu0 == u0_cpu)),
                  "iterations_equal": bool(np.array_equal(git.astype(np.int64), it_cpu.astype(np.int64))),
                  "source": "oracle/_ref (the reference's dlib::mpc)" if kind == "reference" else "oracle/ (C restatement)",
                  "iterations_source": "oracle/liboracle.so (dlib::mpc hides its counter, mpc.h:270)",
                  "what": "EXACT FP64 results of the timed batch vs the CPU leg's solutions of the same instances"}
    # p50 latency of ONE solve on one host core (BASELINE metric: "p50 solve latency"): 96 single-instance calls
    lat = []
    for k in range(96):
        t1 = time.perf_counter()
        run(W.bicycle4_model(), N_, x0[k:k + 1], 1)
        lat.append(time.perf_counter() - t1)
    lat = np.array(lat[16:]) * 1e6
    return parity, {"value": sample / secs, "unit": UNIT, "cores": cores, "kind": kind, "seconds": secs,
            "latency_us_single_solve": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)),
                                        "what": "one cold (4,1,30) solve on one host thread (incl. the ctypes call)"},
            "sample": f"the first {sample} instances of rank 0's batch (seed {seed}), one cold dlib::mpc<4,1,30> per "
                      f"instance, {cores} host threads over contiguous slices, g++ -O2 -ffp-contract=off"}


def run_reference(args):
    """--impl reference: the reference's own CPU solver on this box's host cores; rank 0 only."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    world = max(1, args.gpus)
    kind, run = cpu_solver()
    cores = host_cores()
    sample = args.cpu_sample or 16384
    x0 = W.random_offsets_x0(sample, SEED)
    model = W.bicycle4_model()
    for _ in range(args.warmup):
        run(model, N_, x0[: max(cores * 8, 256)], cores)
    t = 0.0
    for _ in range(args.steps):
        t += run(model, N_, x0, cores)[0]
    val = sample * args.steps / t
    cfg = config_dict(args, world)
    line = {"impl": "reference", "metric": METRIC, "value": val, "unit": UNIT, "n_gpus": args.gpus, "steps": args.steps,
            "warmup": args.warmup, "ms_per_step": 1e3 * t / args.steps, "higher_is_better": True, "scaling": "weak",
            "vs_baseline": None, "dtype": "f64", "data": "synthetic", "config": cfg,
            "cpu_baseline": {"value": val, "unit": UNIT, "cores": cores, "kind": kind,
                             "sample": f"each step = the first {sample} instances of the workload (seed {SEED}) on "
                                       f"{cores} host threads, one cold dlib::mpc<4,1,30> per instance"},
            "e2e": {"value": val, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
            "gpu_launches": 0}
    _emit(line)


# ------------------------------------------------------------------------------------------------- GPU arm
def run_ours(args):
    import torch
    import torch.distributed as dist

    from trajectory_controller_b200.engine import EXACT, FAST, BatchedMPC

    rank = int(os.environ.get("RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    local = int(os.environ.get("LOCAL_RANK", "0"))
    if not torch.cuda.is_available():
        raise SystemExit("bench.py: no CUDA device — the engine has no CPU path (use --impl reference for the CPU arm)")
    torch.cuda.set_device(local)
    if world > 1:
        os.environ.setdefault("MASTER_ADDR", "127.0.0.1")
        os.environ.setdefault("NCCL_DEBUG_FILE", "/dev/stderr")     # stdout carries exactly one JSON line
        dist.init_process_group("nccl", device_id=torch.device("cuda", local))
    dev = torch.device("cuda", local)

    def barrier():
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()

    def max_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    def sum_over_ranks(x: float) -> float:
        if world == 1:
            return x
        t = torch.tensor([x], dtype=torch.float64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.SUM)
        return float(t.item())

    B, K, Wm = args.batch, args.steps, max(args.warmup, 0)
    mode = EXACT if args.mode == "exact" else FAST
    model = W.bicycle4_model()
    x0 = W.random_offsets_x0(B, SEED + rank)
    # pinned host buffers for the end-to-end leg (numpy views of torch pinned tensors)
    px0 = torch.empty((B, S_), dtype=torch.float64).pin_memory(); px0.numpy()[:] = x0
    pu0 = torch.empty((B, I_), dtype=torch.float64).pin_memory()
    pit = torch.empty(B, dtype=torch.int32).pin_memory()
    pst = torch.empty(B, dtype=torch.uint8).pin_memory()
    out = (pu0.numpy(), pit.numpy().view(np.uint32), pst.numpy())
    flush = torch.empty(256 << 20, dtype=torch.uint8, device=dev)
    stream = torch.cuda.Stream(device=dev)

    from trajectory_controller_b200 import engine as _engine
    ver = _engine.lib().bmpc_version().decode()
    build_id = ver.split(" build ")[1].split(" ")[0] if " build " in ver else "unknown"
    eng = BatchedMPC(S_, I_, N_, B, args.precision, mode, device=local)
    eng.set_model(model)
    peak_tf, peak_gi = eng.probe_peak(args.precision, True)
    _, peak_gi_unfused = eng.probe_peak(args.precision, False)
    eng.set_stream(stream.cuda_stream)
    info = eng.kernel_info()

    with torch.cuda.stream(stream):
        # ---------------- resident leg: x0 already in HBM --------------------------------------------------------
        eng.upload_x0(px0.numpy())
        for _ in range(Wm):
            eng.solve_resident_async(B)
        ev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(K)]
        barrier()
        # one sampler per job (rank 0), NVML in-process: nothing is forked beside the host threads that drive the loop
        sampler = ClockSampler(local, period=0.02)
        if rank == 0:
            sampler.start()
        l0 = eng.launch_count()
        for a, b in ev:
            flush.zero_()
            a.record(stream)
            eng.solve_resident_async(B)
            b.record(stream)
        stream.synchronize()
        launches = eng.launch_count() - l0
        barrier()
        step_ms = [a.elapsed_time(b) for a, b in ev]
        t_res = max_over_ranks(sum(step_ms) / 1e3)
        u0_timed, iters, status = eng.download(B, out)
        u0_timed = u0_timed.copy()
        iters = iters.astype(np.int64)
        # per-stage device times of the two-stage solve (events inside libbmpc, three extra untimed solves)
        stage = []
        for _ in range(3):
            eng.solve_resident_async(B); eng.sync()
            stage.append(eng.last_stage_ms())
        stage_a_ms, stage_b_ms = (float(np.mean([x[i] for x in stage])) for i in (0, 1))
        try:
            info_b = eng.pg_kernel_info() if stage_b_ms > 0 else None
        except Exception:
            info_b = None

        # ---------------- end-to-end leg: host buffers through bmpc_solve ------------------------------------------
        # N > 1: plus the job's only collective, the final result gather: one NCCL all_gather each of
        # u0 / iteration counts / status straight from the engine's result buffers in HBM, then a D2H of the gathered
        # arrays into pinned host memory on every rank — every rank ends the step holding the whole job's results.
        gather = None
        if world > 1:
            # u0 | iters | status of a rank are one device allocation (bmpc_device_results_block): ONE all_gather, one D2H
            blk_view, off_it, off_st = eng.device_results_block()
            dblk = torch.as_tensor(blk_view, device=dev)
            nblk = dblk.numel()
            g_blk = torch.empty((world, nblk), dtype=torch.uint8, device=dev)
            h_blk = torch.empty((world, nblk), dtype=torch.uint8).pin_memory()

            gstream = torch.cuda.Stream(device=dev)
            ev_solved, ev_gathered = torch.cuda.Event(), torch.cuda.Event()

            def gather():
                # The collective runs on a side stream after the solve; the next step may start once the all_gather has
                # read this rank's block (ev_gathered), while the D2H of the gathered blocks overlaps that next solve.
                # Everything is complete before the timed region ends (both streams are synchronised there).
                ev_solved.record(stream); gstream.wait_event(ev_solved)
                with torch.cuda.stream(gstream):
                    dist.all_gather_into_tensor(g_blk.view(-1), dblk)
                    ev_gathered.record(gstream)
                    if rank == 0:                      # every rank holds the job's results in HBM; rank 0 also on the host
                        h_blk.copy_(g_blk, non_blocking=True)
                stream.wait_event(ev_gathered)
        for _ in range(min(Wm, 3)):
            eng.solve(px0.numpy(), out=out)
            if gather:
                gather()
        barrier()
        t0 = time.perf_counter()
        for _ in range(K):
            eng.solve(px0.numpy(), out=out)
            if gather:
                gather()
        stream.synchronize()
        if gather:
            gstream.synchronize()
        t_e2e_local = time.perf_counter() - t0
        barrier()
        t_e2e = max_over_ranks(t_e2e_local)
        clocks = sampler.result() if rank == 0 else None
        gather_info = None
        if gather:
            # the gathered copy must be what each rank computed: rank r's block equals its own results
            gstream.synchronize()
            mine_blk = g_blk[rank].cpu().numpy() if rank != 0 else h_blk.numpy()[rank]
            es = 8 if args.precision == 64 else 4
            g_u = mine_blk[: B * I_ * es].view(np.float64 if es == 8 else np.float32).reshape(I_, B).T.astype(np.float64)
            g_i = mine_blk[off_it: off_it + 4 * B].view(np.uint32)
            ok = bool(np.array_equal(g_u, out[0]) and np.array_equal(g_i, out[1]))
            # the collective alone, ranks aligned by a barrier (otherwise the first one measures the skew between ranks)
            gev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(6)]
            for a, b in gev:
                barrier()
                a.record(stream); gather(); stream.wait_stream(gstream); b.record(stream)
            stream.synchronize()
            gms = max_over_ranks(float(np.median([a.elapsed_time(b) for a, b in gev[1:]])))
            gather_info = {"ms": gms, "bytes_per_rank_out": int(world * nblk),
                           "own_block_matches": bool(max_over_ranks(0.0 if ok else 1.0) == 0.0),
                           "what": "one ncclAllGather of every rank's result block (u0 | iters | status, one allocation) from HBM "
                                   "(afterwards every rank holds the whole job's results in HBM) + one D2H of the gathered blocks into pinned host "
                                   "memory on rank 0 (side stream: the copy overlaps the next step's solve), inside e2e; every rank's own "
                                   "results also reach its host buffers through bmpc_solve"}

        # ---------------- single-solve latency through the ABI (batch of 1, cold) ----------------------------------
        lat = []
        one = (np.empty((1, I_)), np.empty(1, dtype=np.uint32), np.empty(1, dtype=np.uint8))
        for k in range(60):
            t1 = time.perf_counter()
            eng.solve(x0[k % B:k % B + 1], out=one)
            lat.append(time.perf_counter() - t1)
        lat = np.array(lat[10:]) * 1e6

    # ---------------- side measurement: the same batch in FAST mode (FMA allowed), checked against this run -------
    fast_mode = None
    if args.mode == "exact" and not args.no_fast:
        u0_exact = out[0].copy()
        with torch.cuda.stream(stream):
            with BatchedMPC(S_, I_, N_, B, args.precision, FAST, device=local) as fe:
                fe.set_model(model); fe.set_stream(stream.cuda_stream); fe.upload_x0(px0.numpy())
                for _ in range(2):
                    fe.solve_resident_async(B)
                fev = [(torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)) for _ in range(min(K, 5))]
                for a, b in fev:
                    flush.zero_(); a.record(stream); fe.solve_resident_async(B); b.record(stream)
                stream.synchronize()
                fu0, fit, _ = fe.download(B)
        fms = float(np.mean([a.elapsed_time(b) for a, b in fev]))
        tol = 1e-6 if args.precision == 64 else 1e-3
        du = np.abs(fu0 - u0_exact)
        fast_mode = {"value_per_gpu": B / fms * 1e3, "unit": UNIT, "ms_per_step": fms, "max_abs_du0_vs_exact": float(du.max()),
                     "tolerance": tol, "within_tolerance": bool(du.max() <= tol),
                     "iteration_counts_equal_frac": float((fit.astype(np.int64) == iters).mean()),
                     "what": "same batch, FMA contraction allowed (BMPC_MODE_FAST); north-star tolerance |du| <= 1e-6 (FP64) / 1e-3 (FP32)"}

    flops = float(W.flops_setup(S_, I_, N_) * B + W.flops_per_iteration(S_, I_, N_) * iters.sum())
    ms_launch = float(np.mean(step_ms))
    achieved_solve = flops / (ms_launch * 1e-3) / 1e12
    tname = "double" if args.precision == 64 else "float"
    mname = "EXACT" if args.mode == "exact" else "FAST"
    if stage_b_ms > 0:
        # dominant kernel = stage B (projected-gradient phase): its share of the step from the in-library events,
        # its algorithmic flops = F_iter x the iterations beyond the 50 coordinate steps
        share_b = stage_b_ms / (stage_a_ms + stage_b_ms)
        kernel_ms = ms_launch * share_b
        kflops = float(W.flops_per_iteration(S_, I_, N_) * np.maximum(iters - 50, 0).sum())
        kname = (f"pg_kernel<{tname},{S_},{I_},{N_},{mname},shared-model> (stage B, one instance per lane) + "
                 f"pair_kernel<{tname},{N_},{mname}> (stage C, its last instances on lane pairs): the projected-gradient phase")
        klaunch = info_b
    else:
        share_b, kernel_ms, kflops, klaunch = 0.0, ms_launch, flops, info
        kname = f"solve_kernel<{tname},{S_},{I_},{N_},{mname},shared-model>"
    achieved = kflops / (kernel_ms * 1e-3) / 1e12
    total_units = B * world
    value = total_units * K / t_res
    e2e = total_units * K / t_e2e
    mean_iters = sum_over_ranks(float(iters.mean())) / world
    eng.close()
    if world > 1:
        dist.destroy_process_group()
    if rank != 0:
        return
    traffic, traffic_src = (None, None)
    if args.precision == 64 and args.mode == "exact" and B == BATCH:
        traffic, traffic_src = ncu_traffic("pg_kernel" if stage_b_ms > 0 else "solve_kernel", build_id)
        if stage_b_ms > 0 and traffic is not None:                      # stage C belongs to the same phase
            t2, _ = ncu_traffic("pair_kernel", build_id)
            traffic += t2 or 0.0
    line = {"metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": K, "warmup": Wm,
            "ms_per_step": 1e3 * t_res / K, "higher_is_better": True, "scaling": "weak", "vs_baseline": None,
            "dtype": "f64" if args.precision == 64 else "f32", "data": "synthetic", "config": config_dict(args, world),
            "clocks": clocks,
            "e2e": {"value": e2e, "unit": UNIT, "ms_per_step": 1e3 * t_e2e / K, "h2d_bytes_per_step": world * B * S_ * 8,
                    "d2h_bytes_per_step": world * B * (I_ * 8 + 4 + 1) + (world * int(nblk) if world > 1 else 0),
                    "api": "bmpc_solve (C ABI, pinned host buffers)" + (" + final result gather (NCCL all_gather + D2H on every rank)" if world > 1 else ""),
                    "gather": gather_info},
            "gpu_launches": int(launches),
            "roofline": {"bound": "fp64_pipe" if args.precision == 64 else "fp32_pipe",
                         "bound_note": "neither 'hbm' nor 'tensor': the solver is chains of small dependent mat-vecs held on chip "
                                       "(DRAM traffic < 0.2 % of peak, no dense contraction), bound by the CUDA-core FP pipe "
                                       "(DESIGN.md 4.1, SURVEY.md 8(d))",
                         "achieved": achieved, "peak": peak_tf,
                         "unit": "TFLOP/s", "frac": achieved / peak_tf,
                         "traffic": traffic, "traffic_unit": "bytes of DRAM traffic per launch (ncu --set full of this build: " + str(traffic_src) + ")",
                         "build_id": build_id,
                         "peak_source": "bmpc_probe_peak: register-resident FMA chains on this device (MEASURED_PEAKS.json "
                                        "holds only HBM and bf16-tensor figures; the solver uses neither)",
                         "peak_ginstr_fused": peak_gi, "peak_ginstr_unfused": peak_gi_unfused,
                         "exact_mode_ceiling_frac": 0.5 if args.mode == "exact" else 1.0,
                         "kernel": kname, "kernel_ms": kernel_ms, "kernel_share_of_step": share_b if stage_b_ms > 0 else 1.0,
                         "algorithmic_flops_per_launch": kflops,
                         # what the hardware executes for one algorithmic unit (one iteration of one instance): the
                         # pattern-specialised EXACT kernel drops the structural zeros / ones of the bicycle model, so it
                         # executes fewer flops than the dense count `achieved` is defined on, and it
                         # may not fuse (1 flop per instruction) — profiles/r2_sass_mix_pg_kernel.txt, pass loop only
                         "executed": {"fp64_pipe_instructions_per_unit": 1033 if (args.precision == 64 and args.mode == "exact") else None,
                                      "flops_per_unit": 943 if (args.precision == 64 and args.mode == "exact") else None,
                                      "algorithmic_flops_per_unit": int(W.flops_per_iteration(S_, I_, N_)),
                                      "fp64_pipe_instr_per_s": (1033.0 * float(np.maximum(iters - 50, 0).sum()) / (kernel_ms * 1e-3) / 1e9)
                                      if (args.precision == 64 and args.mode == "exact" and stage_b_ms > 0) else None,
                                      "unit": "G thread-instructions/s (compare peak_ginstr_unfused)",
                                      "source": "SASS of pg_kernel<double,4,1,30,EXACT,bicycle4,SYM>: 488 DMUL + 455 DADD + 90 DSETP per pass"},
                         "hbm_bytes_per_launch_algorithmic": B * (S_ * 8 * 2 + I_ * 8 + 5) + (B * 3 * N_ * I_ * 8 * 2 if stage_b_ms > 0 else 0),
                         "launch": klaunch,
                         "whole_solve": {"achieved": achieved_solve, "frac": achieved_solve / peak_tf,
                                         "algorithmic_flops_per_step": flops, "ms_per_step": ms_launch,
                                         "stage_a_ms": stage_a_ms, "stage_b_ms": stage_b_ms,
                                         "stage_a_kernel": f"solve_kernel<{tname},{S_},{I_},{N_},{mname},shared-model>",
                                         "stage_a_launch": info}},
            "iterations": {"mean": mean_iters, "p50": float(np.percentile(iters, 50)), "p99": float(np.percentile(iters, 99)),
                           "max": int(iters.max()), "hit_max_iterations": int((status != 0).sum())},
            "latency_us_batch1": {"p50": float(np.percentile(lat, 50)), "p99": float(np.percentile(lat, 99)),
                                  "what": "one cold (4,1,30) solve through bmpc_solve incl. copies"}}
    if fast_mode is not None:
        line["fast_mode"] = fast_mode
    if not args.no_cpu and world == 1:                    # reported baseline: rank 0 at N = 1 only
        check = (u0_timed, iters) if (args.mode == "exact" and args.precision == 64) else None
        parity, line["cpu_baseline"] = cpu_baseline(args.cpu_sample or min(B, 32768), SEED, check)
        if parity is not None:
            line["parity"] = parity
    _emit(line)


def main():
    args = parse_args()
    # stdout must carry exactly one JSON line: libraries (NCCL's version banner, torchrun notices) write to fd 1
    # directly, so fd 1 is pointed at stderr for the whole run and the JSON line goes to the saved descriptor.
    sys.stdout.flush()
    saved = os.dup(1)
    os.dup2(2, 1)
    global _emit
    def _emit(line: dict):
        os.write(saved, (json.dumps(line) + "\n").encode())
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
