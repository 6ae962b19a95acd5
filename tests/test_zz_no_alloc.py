"""b200_dot_noalloc (include/b200_belos.h): the allocation-free variant of MvDot, through the C-ABI.

Collected last on purpose.  Checked against the existing entry point
b200_mv_dot on the same device data, two calls in a row (the static scratch array keeps the first call's
contents), with the tolerance of the MvDot parity test (tests/test_gpu_kernels.py test_dot_norm).
"""
import ctypes as C

import numpy as np
import pytest

import trilinos_b200 as tb
from trilinos_b200._lib import B200Error, check, lib

TOL = 1e-13
MAX_COLS = 32  # B200_NOALLOC_MAX_COLS


def _dev_ptr(mv, col=0):
    p, ld = C.c_void_p(), C.c_int64()
    check(lib.b200_mv_device_ptr(mv._h, col, C.byref(p), C.byref(ld)))
    return p, ld.value


def _noalloc_dot(ctx, A, B, out_mv, n, k):
    (pa, lda), (pb, ldb), (po, _) = _dev_ptr(A), _dev_ptr(B), _dev_ptr(out_mv)
    check(lib.b200_dot_noalloc(pa, lda, pb, ldb, n, k, po, ctx.stream))
    return out_mv.to_numpy()[:k, 0].copy()  # (the copy synchronises on the same stream)


@pytest.mark.gpu
@pytest.mark.parametrize("n,k", [(1, 1), (33, 5), (100003, 4), (1 << 20, 8), (1_000_003, MAX_COLS)])
def test_noalloc_dot_matches_mv_dot_twice_in_a_row(ctx, n, k):
    rng = np.random.default_rng(100 + k)
    out = tb.MultiVec(ctx, MAX_COLS, 1)
    for call in range(2):  # second call: other data, other grid contents left behind in the scratch array
        a = np.asfortranarray(rng.uniform(-1, 1, size=(n, k)))
        b = np.asfortranarray(rng.uniform(-1, 1, size=(n, k)))
        A, B = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, b)
        ref_ab, ref_aa, ref_bb = A.MvDot(B), A.MvDot(A), B.MvDot(B)
        got_ab = _noalloc_dot(ctx, A, B, out, n, k)
        got_aa = _noalloc_dot(ctx, A, A, out, n, k)
        # a.b of independent vectors cancels: bound relative to |a||b|; a.a has no cancellation
        assert np.abs(got_ab - ref_ab).max() <= TOL * np.sqrt(ref_aa * ref_bb).max(), (call, n, k)
        assert np.abs(got_aa - ref_aa).max() <= TOL * np.abs(ref_aa).max(), (call, n, k)
    # a smaller problem right after a large one: stale partials of the larger grid must not leak in
    a = np.asfortranarray(rng.uniform(-1, 1, size=(7, 2)))
    A = tb.MultiVec.from_numpy(ctx, a)
    got = _noalloc_dot(ctx, A, A, out, 7, 2)
    assert np.abs(got - A.MvDot(A)).max() <= TOL * np.abs(got).max()


@pytest.mark.gpu
def test_noalloc_dot_rejects_sizes_beyond_the_static_array(ctx):
    A = tb.MultiVec(ctx, 64, MAX_COLS + 1)
    out = tb.MultiVec(ctx, MAX_COLS + 1, 1)
    (pa, lda), (po, _) = _dev_ptr(A), _dev_ptr(out)
    with pytest.raises(B200Error, match="at most 32"):
        check(lib.b200_dot_noalloc(pa, lda, pa, lda, 64, MAX_COLS + 1, po, ctx.stream))
    with pytest.raises(B200Error):
        check(lib.b200_dot_noalloc(pa, lda, pa, lda, -1, 1, po, ctx.stream))
    check(lib.b200_dot_noalloc(pa, lda, pa, lda, 64, 0, po, ctx.stream))  # nothing to do is not an error


def test_noalloc_dot_size_check_needs_no_device():
    """The limit is checked before anything touches CUDA: same answer on a box without a GPU."""
    buf = (C.c_double * 4)()
    p = C.cast(buf, C.c_void_p)
    with pytest.raises(B200Error, match="at most 32"):
        check(lib.b200_dot_noalloc(p, 4, p, 4, 4, MAX_COLS + 1, p, None))
    assert lib.b200_dot_noalloc(p, 4, p, 4, 4, 0, p, None) == 0
