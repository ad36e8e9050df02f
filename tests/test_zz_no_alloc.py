"""GPU test of snk_match_knn2_noalloc, the kNN-2 entry point for callers that may not allocate device memory
(see include/snakeorb.h). It is called through the C ABI and must return exactly what snk_match_knn2 returns
(bit-equal records — the tolerance of every matcher test here is zero) and what the CPU oracle's matchKnn2 returns
(the two entry points share a kernel, so agreeing with each other alone would not catch a common fault), on two
calls in a row: the calls share one
fixed scratch array inside the library, so the second, smaller call runs on top of the first call's leftovers.

The file name sorts last on purpose: with `pytest -x` every other test has already run when this one does."""
import ctypes as C

import numpy as np
import pytest

pytestmark = pytest.mark.gpu


def _knn2_noalloc(so, d1, d2, device=0, stream=None):
    d1 = np.ascontiguousarray(d1, np.uint64).reshape(-1, 4)
    d2 = np.ascontiguousarray(d2, np.uint64).reshape(-1, 4)
    out = np.zeros(max(len(d1), 1), so.KNN2)
    rc = so.lib().snk_match_knn2_noalloc(device, C.c_void_p(stream), C.c_void_p(d1.ctypes.data), len(d1),
                                         C.c_void_p(d2.ctypes.data), len(d2), C.c_void_p(out.ctypes.data))
    return rc, out[:len(d1)]


def test_knn2_noalloc_equals_knn2_on_two_calls_in_a_row():
    import oracle_py as oracle
    import snakeorb as so

    oracle.build()

    rng = np.random.default_rng(2024)
    ctx = so.Context()
    base = rng.integers(0, 2**63, (64, 4)).astype(np.uint64)
    # call 1: large, tie-heavy (train set = noisy copies of few descriptors); call 2: smaller, different data
    problems = [
        (rng.integers(0, 2**63, (3000, 4)).astype(np.uint64), base[rng.integers(0, 64, 4096)]),
        (base[rng.integers(0, 64, 700)] ^ np.uint64(5), rng.integers(0, 2**63, (130, 4)).astype(np.uint64)),
    ]
    for d1, d2 in problems:
        rc, got = _knn2_noalloc(so, d1, d2)
        assert rc == 0, so.lib().snk_last_error()
        want = ctx.matchKnn2(d1, d2)  # the existing, allocating entry point
        assert got.tobytes() == want.tobytes()
        assert got.tobytes() == oracle.match_knn2(d1, d2).tobytes()  # BruteForceMatcher::matchKnn2 on the CPU
    # empty train set and single query behave like the existing entry point too
    rc, got = _knn2_noalloc(so, problems[0][0][:1], problems[0][1][:0])
    assert rc == 0 and got.tobytes() == ctx.matchKnn2(problems[0][0][:1], problems[0][1][:0]).tobytes()


def test_knn2_noalloc_rejects_sizes_beyond_its_fixed_scratch():
    import snakeorb as so

    big = np.zeros((4097, 4), np.uint64)
    small = np.zeros((8, 4), np.uint64)
    for d1, d2 in ((big, small), (small, big)):
        rc, _ = _knn2_noalloc(so, d1, d2)
        assert rc == -1  # SNK_ERR_INVALID
        assert b"4096" in so.lib().snk_last_error()
