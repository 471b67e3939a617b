"""Multi-GPU host logic on CPU: instance sharding + final gather with world_size 2 over gloo.

The per-rank solver here is the CPU oracle standing in for a GPU rank's BatchedMPC.solve — the thing under test is
the sharding/gather plumbing (trajectory_controller_b200/sharding.py), which is backend-agnostic."""
import os
import socket
import sys

import numpy as np
import pytest
import torch.distributed as dist
import torch.multiprocessing as mp

HERE = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, HERE)

from trajectory_controller_b200 import sharding as SH  # noqa: E402
from trajectory_controller_b200 import workloads as W  # noqa: E402


def test_shard_bounds_cover_and_balance():
    for count in (0, 1, 2, 5, 7, 64, 65536, 65537):
        for world in (1, 2, 3, 4, 8):
            b = SH.shard_bounds(count, world)
            assert b[0][0] == 0 and b[-1][1] == count
            assert all(b[i][1] == b[i + 1][0] for i in range(world - 1))
            sizes = [y - x for x, y in b]
            assert max(sizes) - min(sizes) <= 1
            il = np.concatenate([SH.interleaved_indices(count, r, world) for r in range(world)])
            assert np.array_equal(np.sort(il), np.arange(count))
    with pytest.raises(ValueError):
        SH.shard_bounds(4, 0)


def _free_port():
    s = socket.socket(); s.bind(("127.0.0.1", 0)); p = s.getsockname()[1]; s.close(); return p


def _worker(rank, world, port, count, interleaved, q):
    os.environ["MASTER_ADDR"] = "127.0.0.1"; os.environ["MASTER_PORT"] = str(port)
    dist.init_process_group("gloo", rank=rank, world_size=world)
    try:
        from _oracle import Oracle
        orc = Oracle()
        bike = W.bicycle4_model()
        x0 = W.random_offsets_x0(count, 7)
        seen = []

        def solve(x):
            seen.append(len(x))
            u0, _, it, _ = orc.batch_cold(bike, 10, x, want_controls=False)
            return u0, it, (it >= 10000).astype(np.uint8)

        u0, it, st = SH.solve_sharded(solve, x0, interleaved=interleaved)
        q.put((rank, u0, it, st, seen))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("count,interleaved", [(33, False), (33, True), (1, False), (64, True)])
def test_two_rank_gloo_matches_single_rank(count, interleaved):
    from _oracle import Oracle
    world = 2
    ctx = mp.get_context("spawn")
    q = ctx.Queue()
    port = _free_port()
    procs = [ctx.Process(target=_worker, args=(r, world, port, count, interleaved, q)) for r in range(world)]
    for p in procs:
        p.start()
    got = [q.get(timeout=120) for _ in range(world)]
    for p in procs:
        p.join(timeout=60)
        assert p.exitcode == 0
    x0 = W.random_offsets_x0(count, 7)
    u_ref, _, it_ref, _ = Oracle().batch_cold(W.bicycle4_model(), 10, x0, want_controls=False)
    for rank, u0, it, st, seen in got:
        assert u0.dtype == np.float64 and it.dtype == np.uint32 and st.dtype == np.uint8
        assert np.array_equal(u0.view(np.uint64), u_ref.view(np.uint64)), rank     # every rank holds the full result
        assert np.array_equal(it, it_ref)
        assert st.shape == (count,)
    # work was split, not replicated (a rank with an empty shard only ran the 1-row shape probe)
    work = sorted(s[4][0] for s in got)
    assert sum(work) == count or (count < world and work == [1, 1])
