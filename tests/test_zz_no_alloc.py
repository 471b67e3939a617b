"""bmpc_extract_points_static — the entry point that allocates nothing (its device staging is a fixed array inside
libbmpc.so, include/bmpc.h) — against bmpc_extract_points on the same inputs, through the C ABI, bit for bit, on two
calls in a row (the array keeps the previous call's contents), and the size limit.

Collected last on purpose; it must pass alone in a fresh process and after every other
test.  No counterpart in the reference."""
import numpy as np
import pytest

from trajectory_controller_b200 import engine as E
from trajectory_controller_b200.engine import BatchedMPC, BmpcError, extract_points_static

gpu = pytest.mark.gpu


def _polylines(rng, B, P):
    px = np.cumsum(rng.uniform(0.01, 0.2, (B, P)), axis=1).astype(np.float32)
    py = (0.3 * np.sin(px) + 0.01 * rng.standard_normal((B, P))).astype(np.float32)
    dx, dy, vel = (rng.standard_normal((B, P)).astype(np.float32) for _ in range(3))
    return px, py, dx, dy, vel


def _same(got, want, what):
    assert np.array_equal(got[0], want[0]), what          # selected index
    assert np.array_equal(got[1], want[1]), what          # BMPC_PT_* flag
    for g, w in zip(got[2], want[2]):                     # bit-exact floats: the tolerance of the extraction tests
        assert np.array_equal(g.view(np.uint32), w.view(np.uint32)), what


@gpu
def test_static_extract_equals_extract_points_twice_in_a_row():
    rng = np.random.default_rng(77)
    with BatchedMPC(2, 2, 4, 8) as eng:
        # call 1: ragged per-instance polylines, both layouts; call 2 right behind it with a SMALLER, different problem,
        # so that whatever call 1 left in the array lies under and behind call 2's windows
        for call, (B, P) in enumerate(((3000, 48), (1111, 20))):
            arrs = _polylines(rng, B, P)
            dist = rng.uniform(-0.1, 6.0, B).astype(np.float32)
            npts = rng.integers(0, P + 1, B).astype(np.int32); npts[:4] = [0, 1, 2, P]
            for layout, a in (("bp", arrs), ("pb", tuple(np.ascontiguousarray(x.T) for x in arrs))):
                for n in (None, npts):
                    want = eng.extract_points(*a, dist, npoints=n, layout=layout)
                    got = extract_points_static(*a, dist, npoints=n, layout=layout)
                    _same(got, want, (call, layout, n is not None))
        # one shared polyline, many look-ahead distances incl. exact hits of the cumulative length; then the same call
        # AGAIN: identical results from an array that now holds its own previous output
        P = 64
        sx = np.arange(P, dtype=np.float32) * np.float32(0.125); z = np.zeros(P, dtype=np.float32)
        d = np.concatenate([np.arange(0, 70, dtype=np.float32) * np.float32(0.125), rng.uniform(-1, 9, 500).astype(np.float32)])
        want = eng.extract_points(sx, z, z + 1, z, z + 2, d)
        first = extract_points_static(sx, z, z + 1, z, z + 2, d)
        second = extract_points_static(sx, z, z + 1, z, z + 2, d)
        _same(first, want, "shared, first call"); _same(second, want, "shared, second call")
        # a single polyline and the empty trajectory (default point, idx -1)
        one = [a[0] for a in _polylines(rng, 1, 33)]
        _same(extract_points_static(*one, np.float32([1.3])), eng.extract_points(*one, np.float32([1.3])), "single")
        gi, gf, go = extract_points_static(z[:0], z[:0], z[:0], z[:0], z[:0], np.float32([0.7]))
        assert gi[0] == -1 and gf[0] == 2 and go[0][0] == np.float32(0.7) and go[2][0] == 1.0


@gpu
def test_static_extract_on_a_caller_stream_and_without_a_handle(oracle):
    """No bmpc handle exists here: the entry point needs none.  Results on a caller's stream equal the oracle's."""
    import torch
    rng = np.random.default_rng(78)
    arrs = _polylines(rng, 500, 25)
    dist = rng.uniform(0.0, 3.0, 500).astype(np.float32)
    want = oracle.extract(*arrs, dist, layout="bp")
    st = torch.cuda.Stream(device=0)
    for _ in range(2):
        _same(extract_points_static(*arrs, dist, layout="bp", device=0, stream=st.cuda_stream), want, "stream")


def test_static_extract_limit_matches_the_header():
    import os, re
    hdr = open(os.path.join(os.path.dirname(os.path.dirname(os.path.abspath(__file__))), "include", "bmpc.h")).read()
    m = re.search(r"#define\s+BMPC_STATIC_EXTRACT_BYTES\s+\((\d+)u\s*<<\s*(\d+)\)", hdr)
    assert m and (int(m.group(1)) << int(m.group(2))) == E.STATIC_EXTRACT_BYTES


def test_static_extract_rejects_what_does_not_fit():
    """The size and argument checks come before the first CUDA call: this part runs without a GPU."""
    z = np.zeros(8, dtype=np.float32)
    B = E.STATIC_EXTRACT_BYTES // 4 // 4                  # dist + seven outputs of B elements: about twice the array
    with pytest.raises(BmpcError) as err:
        extract_points_static(z, z, z, z, z, np.zeros(B, dtype=np.float32))
    assert "BMPC_STATIC_EXTRACT_BYTES" in str(err.value) and str(E.STATIC_EXTRACT_BYTES) in str(err.value)
    big = np.zeros((600, 400), dtype=np.float32)          # five arrays of 240,000 floats: 4.8 MB of staging
    with pytest.raises(BmpcError):
        extract_points_static(big, big, big, big, big, np.zeros(600, dtype=np.float32), layout="bp")
    with pytest.raises(BmpcError):                        # and the argument checks of the handle entry point
        extract_points_static(z, z, z, z, z, np.zeros(0, dtype=np.float32))
    with pytest.raises(BmpcError):
        extract_points_static(z, z, z, z, z, np.zeros(4, dtype=np.float32), device=-1)


@gpu
def test_static_extract_still_works_after_a_rejection():
    z = np.zeros(8, dtype=np.float32)
    with pytest.raises(BmpcError):
        extract_points_static(z, z, z, z, z, np.zeros(E.STATIC_EXTRACT_BYTES // 4, dtype=np.float32))
    gi, gf, _ = extract_points_static(np.arange(8, dtype=np.float32), z, z + 1, z, z, np.float32([2.5]))
    assert gi[0] == 3 and gf[0] == 0
