"""Device groups (bmpc_group_*, include/bmpc.h): one controller population sharded by instance over several GPUs of
one process must give, bit for bit, what one handle gives for the same instances.   [-m gpu]

With one GPU visible the group is built over devices [0, 0, 0] (three members sharing the device): the sharding, packing,
worker threads and gather are the same code; with more GPUs it spans them."""
import numpy as np
import pytest

from _oracle import same_values
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import (EXACT, FP64, GROUP_CONTIGUOUS, GROUP_INTERLEAVED, BatchedMPC, BmpcError,
                                               ShardedMPC)

pytestmark = pytest.mark.gpu


def _devices():
    import torch
    n = torch.cuda.device_count()
    return list(range(n)) if n >= 2 else [0, 0, 0]


@pytest.mark.parametrize("assignment", [GROUP_INTERLEAVED, GROUP_CONTIGUOUS])
def test_group_cold_solve_equals_single_handle(oracle, assignment):
    B, N = 10007, 30                       # a prime: ragged shards
    bike = W.bicycle4_model()
    x0 = W.random_offsets_x0(B, 77)
    with BatchedMPC(4, 1, N, B) as one:
        one.set_model(bike)
        u1, it1, st1 = one.solve(x0, keep_controls=True)
        c1 = one.get_controls(B)
    devs = _devices()
    with ShardedMPC(4, 1, N, B, FP64, EXACT, devices=devs, assignment=assignment) as grp:
        assert grp.size() == len(devs)
        assert sum(grp.shard(k)[1] for k in range(grp.size())) == B
        grp.set_model(bike)
        ug, itg, stg = grp.solve(x0, keep_controls=True)
        assert np.array_equal(ug.view(np.uint64), u1.view(np.uint64))
        assert np.array_equal(itg, it1) and np.array_equal(stg, st1)
        assert np.array_equal(grp.get_controls(B).view(np.uint64), c1.view(np.uint64))
        part = grp.get_controls(1000, first=4321)
        assert np.array_equal(part.view(np.uint64), c1[4321:5321].view(np.uint64))
        # a prefix of the population (count < capacity)
        up, itp, _ = grp.solve(x0[:1234])
        assert np.array_equal(up.view(np.uint64), u1[:1234].view(np.uint64)) and np.array_equal(itp, it1[:1234])
        assert grp.last_kernel_ms() > 0
    pick = np.arange(0, B, 97)
    uo, _, ito, _ = oracle.batch_cold(bike, N, x0[pick], nthreads=4, want_controls=False)
    assert same_values(ug[pick], uo) and np.array_equal(itg[pick], ito)


@pytest.mark.parametrize("assignment", [GROUP_INTERLEAVED, GROUP_CONTIGUOUS])
def test_group_per_instance_models_and_warm_sequence(assignment):
    """Blocked parameter sweep (per-instance models), per-instance targets, persistent controllers over 4 ticks."""
    models, x0 = W.sweep_models(10, 16, 44)          # 64 combos x 16 x0
    B = len(x0)
    rng = np.random.default_rng(5)
    tgt = np.zeros((B, 4)); tgt[:, 0] = 0.05 * rng.standard_normal(B)
    devs = _devices()
    with BatchedMPC(4, 1, 10, B) as one, ShardedMPC(4, 1, 10, B, devices=devs, assignment=assignment) as grp:
        for e in (one, grp):
            e.set_model(models, per_instance=True)
            e.set_options(0.001, 5000)
            e.set_target(tgt)
        x = x0.copy()
        for tick in range(4):
            u1, it1, st1 = one.solve(x, warm=True)
            ug, itg, stg = grp.solve(x, warm=True)
            assert np.array_equal(ug.view(np.uint64), u1.view(np.uint64)), tick
            assert np.array_equal(itg, it1) and np.array_equal(stg, st1), tick
            x = x * 0.9 + 0.01 * u1.repeat(4, axis=1)
        assert np.array_equal(grp.get_controls(B).view(np.uint64), one.get_controls(B).view(np.uint64))
        grp.reset(); one.reset()
        assert not grp.get_controls(B).any()


def test_group_errors():
    with pytest.raises(BmpcError):
        ShardedMPC(4, 1, 30, 2, devices=[0, 0, 0])           # capacity < members
    with pytest.raises(BmpcError):
        ShardedMPC(4, 1, 30, 64, devices=[0, 99])            # no such device
    with ShardedMPC(4, 1, 30, 64, devices=[0, 0]) as grp:
        with pytest.raises(BmpcError) as e:
            grp.solve(np.zeros((8, 4)))                      # no model yet: the member's error comes through
        assert "member" in str(e.value) and "set_model" in str(e.value)
