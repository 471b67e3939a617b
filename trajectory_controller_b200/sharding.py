"""Instance sharding of a batch of MPC controllers over the GPUs of one box.

Every controller is independent (no shared mutable state, no exchange step: mpc.h:216-347 touches only its own
object), so the batch is cut along the instance axis, one process per GPU, and the data path has NO collective.
The only communication is the final result gather (u0, iteration counts, status), a single all_gather of a few
bytes per instance; it runs on whatever backend the process group has (NCCL over NVLink on the GPU box, gloo in the
CPU tests).

This module is host logic only: it never touches oracle/, and the solver it drives is whatever callable it is handed
(the engine's BatchedMPC.solve on a GPU rank).
"""
from __future__ import annotations

from typing import Callable, Sequence

import numpy as np


def shard_bounds(count: int, world: int) -> list[tuple[int, int]]:
    """Contiguous, balanced [first, last) slices: the first `count % world` ranks take one extra instance.
    Ragged and empty shards are legal (count < world)."""
    if world < 1 or count < 0:
        raise ValueError("shard_bounds: world >= 1 and count >= 0 required")
    base, extra = divmod(count, world)
    out, first = [], 0
    for r in range(world):
        n = base + (1 if r < extra else 0)
        out.append((first, first + n))
        first += n
    return out


def shard_range(count: int, rank: int, world: int) -> tuple[int, int]:
    return shard_bounds(count, world)[rank]


def interleaved_indices(count: int, rank: int, world: int) -> np.ndarray:
    """Round-robin assignment {i : i mod world == rank}: evens out heavy-tailed iteration counts when neighbouring
    instances are correlated (parameter sweeps sorted by weight)."""
    return np.arange(rank, count, world, dtype=np.int64)


def gather_results(local: Sequence[np.ndarray], count: int, group=None, interleaved: bool = False):
    """All-gather per-shard result arrays (first axis = local instance) into global instance order on every rank.

    `local` arrays must have the same dtypes/trailing shapes on all ranks.  Works with world_size 1 without an
    initialised process group."""
    import torch
    import torch.distributed as dist

    if not (dist.is_available() and dist.is_initialized()):
        return [np.asarray(a) for a in local]
    world, rank = dist.get_world_size(group), dist.get_rank(group)
    backend = dist.get_backend(group)
    dev = torch.device("cuda", torch.cuda.current_device()) if backend == "nccl" else torch.device("cpu")
    bounds = shard_bounds(count, world)
    sizes = [len(interleaved_indices(count, r, world)) if interleaved else b[1] - b[0] for r, b in enumerate(bounds)]
    pad = max(sizes) if sizes else 0
    outs = []
    for a in local:
        a = np.ascontiguousarray(a)
        if a.shape[0] != sizes[rank]:
            raise ValueError(f"gather_results: rank {rank} holds {a.shape[0]} rows, its shard has {sizes[rank]}")
        view = a.view(np.int32) if a.dtype == np.uint32 else a      # torch has no uint32 collectives
        buf = torch.zeros((pad,) + view.shape[1:], dtype=torch.from_numpy(view[:0]).dtype, device=dev)
        if sizes[rank]:
            buf[: sizes[rank]] = torch.from_numpy(view).to(dev)
        parts = [torch.empty_like(buf) for _ in range(world)]
        dist.all_gather(parts, buf, group=group)
        full = np.empty((count,) + a.shape[1:], dtype=a.dtype)
        for r in range(world):
            got = parts[r][: sizes[r]].cpu().numpy()
            if a.dtype == np.uint32:
                got = got.view(np.uint32)
            if interleaved:
                full[interleaved_indices(count, r, world)] = got
            else:
                full[bounds[r][0]: bounds[r][1]] = got
        outs.append(full)
    return outs


def solve_sharded(solve: Callable[[np.ndarray], Sequence[np.ndarray]], x0: np.ndarray, group=None,
                  interleaved: bool = False):
    """Solve the rank's shard of a global batch x0[count][S] with `solve` (returns arrays with first axis = local
    instance, e.g. BatchedMPC.solve -> (u0, iters, status)), then gather every rank's results in instance order."""
    import torch.distributed as dist

    count = x0.shape[0]
    if dist.is_available() and dist.is_initialized():
        world, rank = dist.get_world_size(group), dist.get_rank(group)
    else:
        world, rank = 1, 0
    if interleaved:
        mine = x0[interleaved_indices(count, rank, world)]
    else:
        a, b = shard_range(count, rank, world)
        mine = x0[a:b]
    if mine.shape[0]:
        local = [np.asarray(r) for r in solve(np.ascontiguousarray(mine))]
    else:
        # an empty shard still has to contribute correctly-typed empty rows; learn the row shapes from one dummy row
        probe = [np.asarray(r) for r in solve(np.ascontiguousarray(x0[:1]))]
        local = [p[:0] for p in probe]
    return gather_results(local, count, group, interleaved)
