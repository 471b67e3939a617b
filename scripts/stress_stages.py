"""Developer stress: staged solve (A -> B -> C with hand-over) against stages A + B only (bmpc_set_split 2) over batch sizes
around every scheduling threshold, stopping options and warm sequences — results must be bit-equal, nothing may hang."""
import sys, time
import numpy as np
sys.path.insert(0, '.')
from trajectory_controller_b200 import workloads as W
from trajectory_controller_b200.engine import BatchedMPC
bike = W.bicycle4_model()
CAP = 160000
x0 = W.random_offsets_x0(CAP, 7)
bad = 0
with BatchedMPC(4, 1, 30, CAP) as a, BatchedMPC(4, 1, 30, CAP) as b:
    a.set_model(bike); b.set_model(bike); b.set_split(2)
    for (eps, mi) in ((0.01, 10000), (0.01, 60), (0.01, 51), (0.001, 3000), (0.05, 10000), (0.01, 700)):
        a.set_options(eps, mi); b.set_options(eps, mi)
        for n in (1, 31, 33, 4097, 9471, 9472, 9473, 9600, 18944, 18945, 37887, 37889, 75777, 151553, 160000):
            t0 = time.time()
            ua, ia, sa = a.solve(x0[:n]); ub, ib, sb = b.solve(x0[:n])
            ok = np.array_equal(ua.view(np.uint64), ub.view(np.uint64)) and np.array_equal(ia, ib) and np.array_equal(sa, sb)
            bad += not ok
            print(f"eps {eps} max_iter {mi} n {n}: {'ok' if ok else 'MISMATCH'} iters mean {ia.mean():.1f} max {ia.max()} hit {int((sa & 1).sum())} {time.time() - t0:.2f}s", flush=True)
    # warm sequences with moving targets across the hand-over
    a.set_options(0.01, 10000); b.set_options(0.01, 10000)
    for n in (5000, 40000):
        a.reset(); b.reset()
        x = x0[:n].copy()
        for tick in range(4):
            ua, ia, _ = a.solve(x, warm=True); ub, ib, _ = b.solve(x, warm=True)
            ok = np.array_equal(ua.view(np.uint64), ub.view(np.uint64)) and np.array_equal(ia, ib)
            bad += not ok
            print(f"warm n {n} tick {tick}: {'ok' if ok else 'MISMATCH'} iters mean {ia.mean():.1f}", flush=True)
            x = x * 0.8
            x[:, 2] += 0.1 * ua[:, 0]
        ca, cb = a.get_controls(n), b.get_controls(n)
        bad += not np.array_equal(ca.view(np.uint64), cb.view(np.uint64))
print("stress done, mismatches:", bad)
sys.exit(1 if bad else 0)
