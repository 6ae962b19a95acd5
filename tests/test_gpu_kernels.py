"""Per-kernel parity: CUDA path (through the C-ABI) vs the CPU oracle on the same seeded inputs.

Bar (BASELINE.json north_star): normwise relative error <= 1e-13 for every FP64 kernel; exact
results where the reference's own MVOPTester demands them (belos/src/BelosMVOPTester.hpp).
"""
import numpy as np
import pytest

import oracle
import trilinos_b200 as tb

pytestmark = pytest.mark.gpu
TOL = 1e-13


def relerr(a, b):
    a, b = np.asarray(a, dtype=np.float64), np.asarray(b, dtype=np.float64)
    d = np.linalg.norm((a - b).ravel())
    s = np.linalg.norm(b.ravel())
    return d / s if s > 0 else d


def rnd(n, k, seed):
    return np.asfortranarray(np.random.default_rng(seed).uniform(-1, 1, size=(n, k)))


SIZES = [1, 3, 31, 32, 33, 127, 1000, 4097, 100003]


@pytest.mark.parametrize("n", SIZES)
@pytest.mark.parametrize("k", [1, 4, 5])
def test_dot_norm(ctx, n, k):
    a, b = rnd(n, k, 1), rnd(n, k, 2)
    A, B = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, b)
    # a.b of two independent vectors cancels: the bound is relative to |a||b| (the condition of the sum), not to |a.b|
    assert np.abs(A.MvDot(B) - oracle.dot(a, b)).max() <= TOL * np.sqrt(oracle.dot(a, a) * oracle.dot(b, b)).max()
    assert relerr(A.MvDot(A), oracle.dot(a, a)) <= TOL
    for t in (tb.NORM_ONE, tb.NORM_TWO, tb.NORM_INF):
        assert relerr(A.MvNorm(t), oracle.norm(a, t)) <= TOL
    assert np.array_equal(A.MvNorm(tb.NORM_INF), np.abs(a).max(axis=0))  # max is exact


@pytest.mark.parametrize("n", [1, 33, 1000, 100003])
@pytest.mark.parametrize("alpha,beta", [(1.0, 0.0), (0.0, 1.0), (0.0, 0.0), (2.5, -0.75), (1.0, 1.0), (-1.0, 1.0)])
def test_add_mv(ctx, n, alpha, beta):
    a, b = rnd(n, 4, 3), rnd(n, 4, 4)
    A, B = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, b)
    Y = A.Clone(4)
    Y.MvAddMv(alpha, A, beta, B)
    ref = oracle.add_mv(alpha, a, beta, b)
    y = Y.to_numpy()
    if alpha == 0.0 or beta == 0.0 or (alpha, beta) == (1.0, 1.0) or (alpha, beta) == (-1.0, 1.0):
        assert np.array_equal(y, ref)  # a single rounding at most: must be bit-exact
    else:
        assert relerr(y, ref) <= TOL


def test_add_mv_does_not_read_dropped_operand(ctx):
    """alpha == 0 / beta == 0 must not propagate NaNs from the other operand
    (Epetra_MultiVector.cpp:1348-1355; contract in BelosMVOPTester.hpp:838-844)."""
    n = 1000
    a = rnd(n, 3, 5)
    bad = np.full((n, 3), np.nan, order="F")
    A, Bad = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, bad)
    Y = A.Clone(3)
    Y.MvAddMv(1.0, A, 0.0, Bad)
    assert np.array_equal(Y.to_numpy(), a)
    Y.MvAddMv(0.0, Bad, 1.0, A)
    assert np.array_equal(Y.to_numpy(), a)


def test_add_mv_aliased(ctx):
    """LinearProblem::computeCurrResVec calls MvAddMv(-1,R,1,B,R) (BelosLinearProblem.hpp:1207)."""
    n = 5003
    r, b = rnd(n, 4, 6), rnd(n, 4, 7)
    R, B = tb.MultiVec.from_numpy(ctx, r), tb.MultiVec.from_numpy(ctx, b)
    R.MvAddMv(-1.0, R, 1.0, B)
    assert np.array_equal(R.to_numpy(), oracle.add_mv(-1.0, r, 1.0, b))


@pytest.mark.parametrize("n", [1, 33, 4097, 100003])
def test_scale_init_random(ctx, n):
    a = rnd(n, 5, 8)
    A = tb.MultiVec.from_numpy(ctx, a)
    A.MvScale(-3.5)
    assert np.array_equal(A.to_numpy(), oracle.scale_cols(-3.5, a))
    s = np.array([1.0, 0.0, -2.0, 0.5, 3.0])
    A.MvScale(s)
    assert np.array_equal(A.to_numpy(), oracle.scale_cols(s, oracle.scale_cols(-3.5, a)))
    A.MvInit(7.25)
    assert np.array_equal(A.to_numpy(), np.full((n, 5), 7.25))
    ctx.set_seed(12345)
    A.MvRandom()
    ref, s1 = oracle.random(12345, n, 5)
    assert np.array_equal(A.to_numpy(), ref)  # integer LCG + one division: bit-exact
    A.MvRandom()  # the stream continues where it stopped
    ref2, _ = oracle.random(s1, n, 5)
    assert np.array_equal(A.to_numpy(), ref2)


@pytest.mark.parametrize("n", [1, 5, 127, 1000, 4097, 100003])
@pytest.mark.parametrize("p,q", [(1, 1), (3, 1), (4, 4), (5, 3), (7, 5), (33, 4), (40, 8), (100, 4), (9, 16), (64, 2), (300, 16), (70, 8), (20, 24), (2, 1), (4, 1), (7, 1), (8, 1), (9, 1), (40, 12), (33, 13)])
def test_trans_mv(ctx, n, p, q):
    a, v = rnd(n, p, 9), rnd(n, q, 10)
    A, V = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, v)
    got = V.MvTransMv(1.5, A)
    ref = oracle.trans_mv(1.5, a, v)
    assert relerr(got, ref) <= TOL
    # deterministic: same launch twice gives the same bits
    assert np.array_equal(got, V.MvTransMv(1.5, A))


def test_trans_mv_contract(ctx):
    """BelosMVOPTester.hpp:792-892: alpha == 0 -> exact zeros; zero operand -> exact zeros;
    B.stride() honoured, entries outside the p x q block untouched."""
    n, p, q = 1000, 5, 3
    a, v = rnd(n, p, 11), rnd(n, q, 12)
    A, V = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, v)
    assert np.array_equal(V.MvTransMv(0.0, A), np.zeros((p, q)))
    Z = V.Clone(q)
    Z.MvInit(0.0)
    assert np.array_equal(Z.MvTransMv(1.0, A), np.zeros((p, q)))
    big = np.full((p + 4, q), -7.0, order="F")
    V.MvTransMv(1.0, A, big, ldb=p + 4)
    assert relerr(big[:p], oracle.trans_mv(1.0, a, v)) <= TOL
    assert np.all(big[p:] == -7.0)


@pytest.mark.parametrize("n", [1, 5, 127, 1000, 4097, 100003])
@pytest.mark.parametrize("p,q", [(1, 1), (3, 1), (4, 4), (5, 3), (7, 5), (33, 4), (40, 8), (300, 4), (9, 16), (70, 16), (37, 24), (2, 1), (4, 1), (7, 1), (9, 1)])
@pytest.mark.parametrize("alpha,beta", [(-1.0, 1.0), (1.0, 0.0), (0.5, 2.0)])
def test_times_mat_add_mv(ctx, n, p, q, alpha, beta):
    a, v, b = rnd(n, p, 13), rnd(n, q, 14), rnd(p, q, 15)
    A, V = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, v)
    if beta == 0.0:  # must not read V
        V = tb.MultiVec.from_numpy(ctx, np.full((n, q), np.nan, order="F"))
    V.MvTimesMatAddMv(alpha, A, b, beta)
    ref = oracle.times_mat_add_mv(alpha, a, b, beta, v)
    assert relerr(V.to_numpy(), ref) <= TOL


def test_views_and_setblock(ctx):
    """index list {0,5,2,2,9}: unordered and repeated (BelosMVOPTester.hpp:197-201)."""
    n = 777
    a = rnd(n, 10, 16)
    A = tb.MultiVec.from_numpy(ctx, a)
    idx = [0, 5, 2, 2, 9]
    C1 = A.CloneCopy(idx)
    assert np.array_equal(C1.to_numpy(), a[:, idx])
    Vw = A.CloneView(idx)
    assert np.array_equal(Vw.to_numpy(), a[:, idx])
    assert Vw.GetNumberVecs() == 5 and Vw.GetGlobalLength() == n
    assert relerr(Vw.MvNorm(), oracle.norm(a[:, idx])) <= TOL
    assert relerr(Vw.MvDot(C1), oracle.dot(a[:, idx], a[:, idx])) <= TOL
    g = rnd(n, 4, 17)
    G = tb.MultiVec.from_numpy(ctx, g)
    got = Vw.MvTransMv(1.0, G)
    assert relerr(got, oracle.trans_mv(1.0, g, a[:, idx])) <= TOL
    # writing through a view changes the parent; copies stay
    W = A.CloneViewNonConst([7, 3])
    W.MvInit(2.0)
    a2 = a.copy()
    a2[:, [7, 3]] = 2.0
    assert np.array_equal(A.to_numpy(), a2)
    assert np.array_equal(C1.to_numpy(), a[:, idx])
    blk = rnd(n, 3, 18)
    A.SetBlock(tb.MultiVec.from_numpy(ctx, blk), [8, 1, 4])
    a2[:, [8, 1, 4]] = blk
    assert np.array_equal(A.to_numpy(), a2)
    # update through non-contiguous A and non-contiguous target
    b = rnd(5, 2, 19)
    T = A.CloneViewNonConst([6, 0])
    T.MvTimesMatAddMv(1.0, C1, b, 1.0)
    ref = oracle.times_mat_add_mv(1.0, a[:, idx], b, 1.0, a2[:, [6, 0]])
    assert relerr(T.to_numpy(), ref) <= TOL


def test_error_behaviour(ctx):
    A = tb.MultiVec(ctx, 100, 3)
    B = tb.MultiVec(ctx, 100, 2)
    C_ = tb.MultiVec(ctx, 90, 3)
    with pytest.raises(tb.B200Error):
        A.MvAddMv(1.0, B, 1.0, B)
    with pytest.raises(tb.B200Error):
        A.MvAddMv(1.0, C_, 1.0, C_)
    with pytest.raises(tb.B200Error):
        A.MvScale([1.0, 2.0])
    with pytest.raises(tb.B200Error):
        A.CloneView([0, 3])
    with pytest.raises(tb.B200Error):
        A.MvTimesMatAddMv(1.0, B, np.zeros((3, 3)), 0.0)


MATS = [("Laplace2D", (100, 100, 1)), ("Laplace3D", (20, 21, 22)), ("Brick3D", (15, 14, 13)), ("Recirc2D", (64, 50, 1)),
        ("Tridiag", (1000, 1, 1)), ("Laplace3D", (3, 1, 1)), ("Laplace3D", (40, 40, 40))]


@pytest.mark.parametrize("name,dims", MATS)
@pytest.mark.parametrize("k", [1, 2, 4, 5, 8, 16])
def test_apply(ctx, name, dims, k):
    n, rp, ci, va = oracle.galeri_csr(name, *dims)
    op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
    x = rnd(n, k, 20)
    X = tb.MultiVec.from_numpy(ctx, x)
    Y = X.Clone(k)
    Y.MvInit(np.nan)  # result must not depend on the prior contents of Y (MVOPTester.hpp:1454-1600)
    op.Apply(X, Y)
    ref = oracle.spmm(rp, ci.astype(np.int32), va, x)
    assert relerr(Y.to_numpy(), ref) <= TOL
    assert np.array_equal(X.to_numpy(), x)
    Z = X.Clone(k)
    Z.MvInit(0.0)
    op.Apply(Z, Y)
    assert np.array_equal(Y.to_numpy(), np.zeros((n, k)))  # OP*0 == 0 exactly


def test_apply_wide_rows(ctx):
    """rows longer than one shared-memory stage (4096 nonzeros) and empty rows."""
    rng = np.random.default_rng(21)
    n = 600
    import scipy.sparse as sp
    A = sp.random(n, n, density=0.02, random_state=3, format="lil")
    A[5, :] = rng.uniform(-1, 1, n)
    A[300, :] = rng.uniform(-1, 1, n)
    A[7, :] = 0
    A = A.tocsr()
    A.sort_indices()
    A = sp.vstack([A] * 1).tocsr()
    # make one CTA's segment exceed the stage: 256 dense rows
    D = sp.csr_matrix(rng.uniform(-1, 1, (256, n)))
    A = sp.vstack([D, A[256:]]).tocsr()
    A.sort_indices()
    op = tb.CrsOperator(ctx, n, 0, n, A.indptr, A.indices.astype(np.int64), A.data)
    x = rnd(n, 4, 22)
    X = tb.MultiVec.from_numpy(ctx, x)
    Y = X.Clone(4)
    op.Apply(X, Y)
    assert relerr(Y.to_numpy(), oracle.spmm(A.indptr, A.indices, A.data, x)) <= TOL


def test_full_size_properties(ctx):
    """BASELINE config 2 sizes (n = 1e6, b = 4, basis width 200): size-independent properties."""
    n, p, q = 1_000_000, 200, 4
    ctx.set_seed(7)
    Q = tb.MultiVec(ctx, n, p, zero=False)
    Q.MvRandom()
    X = tb.MultiVec(ctx, n, q, zero=False)
    X.MvRandom()
    G = X.MvTransMv(1.0, Q)
    # Gram block agrees with column-wise dots (two different kernels)
    for j in (0, 57, 199):
        Qj = Q.CloneView([j] * q)
        assert relerr(X.MvDot(Qj), G[j]) <= 1e-12
    # linearity: (Q^T (2X)) == 2 (Q^T X) exactly (power-of-two scaling)
    X2 = X.CloneCopy()
    X2.MvScale(2.0)
    assert np.array_equal(X2.MvTransMv(1.0, Q), 2.0 * G)
    # projection: X - Q (Q^T X) / ... : update then Gram against an orthogonal-ish basis shrinks
    n2 = X.MvNorm()
    X.MvTimesMatAddMv(0.0, Q, G, 1.0)  # alpha = 0: X unchanged
    assert np.array_equal(X.MvNorm(), n2)
    # update consistency: X + Q*E(j) picks column j
    E = np.zeros((p, q))
    E[3, 0] = 1.0
    Y = X.CloneCopy()
    Y.MvTimesMatAddMv(1.0, Q, E, 1.0)
    Y.MvAddMv(1.0, Y, -1.0, X)
    q3 = Q.CloneView([3])
    y0 = Y.CloneView([0])
    d = y0.CloneCopy()
    d.MvAddMv(1.0, y0, -1.0, q3)
    assert d.MvNorm()[0] <= 1e-12 * q3.MvNorm()[0]
    assert np.all(Y.CloneView([1, 2, 3]).MvNorm() == 0.0)


# ----------------------------------------------------------------------------------------------
# Fused reductions / remembered column products: same values as the standalone
# reduction kernels (<= 1e-13 vs the oracle), and never a stale value after a column was rewritten.
@pytest.mark.parametrize("n", [1, 3, 33, 515, 4097, 100003])
@pytest.mark.parametrize("p,q", [(1, 1), (3, 1), (7, 4), (40, 8), (5, 3), (40, 16), (7, 19), (2, 1), (6, 1)])
def test_fused_update_dots(ctx, n, p, q):
    a, v, bm = rnd(n, p, 11), rnd(n, q, 12), rnd(p, q, 13) * 0.1
    A, V = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, v)
    Vold = V.CloneCopy()
    h0 = ctx.dot_hits()
    V.MvTimesMatAddMv(-1.0, A, bm, 1.0)
    vnew = oracle.times_mat_add_mv(-1.0, a, bm, 1.0, v)
    d_self, d_old, nrm = V.MvDot(V), V.MvDot(Vold), V.MvNorm(tb.NORM_TWO)
    assert ctx.dot_hits() - h0 == 3 * q, "the three reductions after the update must come from its epilogue"
    assert relerr(V.to_numpy(), vnew) <= TOL
    assert relerr(d_self, oracle.dot(vnew, vnew)) <= TOL
    assert np.abs(d_old - oracle.dot(vnew, v)).max() <= TOL * np.sqrt(oracle.dot(vnew, vnew) * oracle.dot(v, v)).max()
    assert relerr(nrm, oracle.norm(vnew, 1)) <= TOL
    # same numbers with the fusion switched off (standalone kernels)
    ctx.set_fusion(False)
    try:
        assert relerr(V.MvDot(V), d_self) <= TOL and relerr(V.MvNorm(tb.NORM_TWO), nrm) <= TOL
        assert np.abs(V.MvDot(Vold) - d_old).max() <= TOL * np.sqrt(oracle.dot(vnew, vnew) * oracle.dot(v, v)).max()
    finally:
        ctx.set_fusion(True)


def test_remembered_dots_are_invalidated_by_every_write(ctx):
    n, k = 5003, 3
    a, b = rnd(n, k, 21), rnd(n, k, 22)
    A, B = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, b)

    def check():
        got, ref = A.MvDot(A), oracle.dot(A.to_numpy(), A.to_numpy())
        assert relerr(got, ref) <= TOL
        assert relerr(A.MvNorm(tb.NORM_TWO), np.sqrt(ref)) <= TOL

    check()
    h0 = ctx.dot_hits()
    check()
    assert ctx.dot_hits() - h0 == 2 * k            # unchanged columns: served from the memo
    for write in (lambda: A.MvScale(1.5), lambda: A.MvScale([2.0, 0.5, -1.0]), lambda: A.MvAddMv(1.0, A, 0.25, B),
                  lambda: A.MvInit(0.75), lambda: A.MvRandom(), lambda: A.set_numpy(b * 3.0),
                  lambda: A.SetBlock(B, [2, 0, 1]), lambda: A.MvAddMv(0.0, A, 1.0, B),
                  lambda: A.CloneViewNonConst([1]).MvScale(-3.0),
                  lambda: A.MvTimesMatAddMv(0.5, B, rnd(k, k, 23), 0.0)):
        write()
        h0 = ctx.dot_hits()
        check()
    # a copy carries its source's content id: x.x of the copy is known without a launch
    C2 = A.CloneCopy()
    h0 = ctx.dot_hits()
    assert relerr(C2.MvDot(C2), oracle.dot(A.to_numpy(), A.to_numpy())) <= TOL
    assert ctx.dot_hits() - h0 == k
    # writing the copy must not poison the original (and vice versa)
    C2.MvScale(2.0)
    assert relerr(C2.MvDot(C2), 4.0 * oracle.dot(A.to_numpy(), A.to_numpy())) <= TOL
    check()


@pytest.mark.parametrize("name,dims,k", [("Laplace3D", (9, 8, 7), 4), ("Laplace2D", (33, 31, 1), 1), ("Brick3D", (7, 6, 5), 8),
                                         ("Recirc2D", (20, 17, 1), 3)])
def test_fused_apply_dots(ctx, name, dims, k):
    n, rp, ci, va = tb.galeri(name, *dims)
    op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
    x = rnd(n, k, 31)
    X, Y = tb.MultiVec.from_numpy(ctx, x), tb.MultiVec(ctx, n, k)
    h0 = ctx.dot_hits()
    op.Apply(X, Y)
    yref = oracle.spmm(rp, ci.astype(np.int32), va, x)
    d = Y.MvDot(Y)
    assert ctx.dot_hits() - h0 == k
    assert relerr(Y.to_numpy(), yref) <= TOL
    assert relerr(d, oracle.dot(yref, yref)) <= TOL
    op.Apply(X, Y)                                  # rewritten: ids change, values identical
    assert relerr(Y.MvNorm(tb.NORM_TWO), oracle.norm(yref, 1)) <= TOL


# ---- deferred single-column MvScale / lazy single-column CloneCopy (capi.cu prep_read / prep_write) ----------------
def test_findbasis_sequence_without_copy_and_scale_launches(ctx):
    """The per-column call sequence of DGKS findBasis (BelosDGKSOrthoManager.hpp:1000-1225): CloneCopy of column j,
    MvTransMv against the previous columns, MvTimesMatAddMv, the two MvDot, MvScale.  Values must equal the oracle's;
    the copy is filled by the update kernel and the scale by the next skinny Gram -- no launch of their own."""
    n, b = 20011, 4
    x = rnd(n, b, 41)
    X = tb.MultiVec.from_numpy(ctx, x)
    ref = x.copy()
    lz0, sc0 = ctx.deferral_stats()
    l0, _ = ctx.counters()
    for j in range(b):
        Xj = X.CloneViewNonConst([j])
        old = Xj.CloneCopy() if j > 0 else None
        old_dot = Xj.MvDot(Xj)
        assert relerr(old_dot, oracle.dot(ref[:, [j]], ref[:, [j]])) <= TOL
        if j > 0:
            prev = X.CloneView(list(range(j)))
            prod = Xj.MvTransMv(1.0, prev)
            assert relerr(prod, oracle.trans_mv(1.0, ref[:, :j], ref[:, [j]])) <= TOL
            refold = ref[:, [j]].copy()
            Xj.MvTimesMatAddMv(-1.0, prev, prod, 1.0)
            ref[:, [j]] = oracle.times_mat_add_mv(-1.0, ref[:, :j], prod, 1.0, ref[:, [j]])
            new_dot = Xj.MvDot(Xj)
            assert relerr(new_dot, oracle.dot(ref[:, [j]], ref[:, [j]])) <= TOL
            od = Xj.MvDot(old)
            assert np.abs(od - oracle.dot(ref[:, [j]], refold)).max() <= TOL * np.sqrt(new_dot * old_dot).max()
            assert np.array_equal(old.to_numpy(), refold), "the lazy copy must hold the column as it was"
        else:
            new_dot = old_dot
        s = 1.0 / np.sqrt(new_dot[0])
        Xj.MvScale(s)
        ref[:, j] = ref[:, j] * s
    l1, _ = ctx.counters()
    lz1, sc1 = ctx.deferral_stats()
    assert lz1 - lz0 == b - 1, "every CloneCopy of the sequence must have been served by the update kernel"
    assert sc1 - sc0 == b - 1, "the scales of columns 0..b-2 must have been folded into the next skinny Gram"
    assert np.array_equal(X.to_numpy(), ref) or relerr(X.to_numpy(), ref) <= TOL   # (reading flushes the last scale)
    q = oracle.trans_mv(1.0, X.to_numpy(), X.to_numpy())
    assert np.abs(q - np.eye(b)).max() <= 1e-12


def test_deferred_scale_is_visible_to_every_reader(ctx):
    n = 7001
    a = rnd(n, 3, 51)
    for reader in ("to_numpy", "norm", "dot", "addmv", "apply_view", "gram_block", "clonecopy", "setblock", "update_src",
                   "scale_again", "device_overwrite"):
        A = tb.MultiVec.from_numpy(ctx, a)
        c1 = A.CloneViewNonConst([1])
        c1.MvScale(-2.5)                       # deferred
        want = a.copy()
        want[:, 1] *= -2.5
        if reader == "to_numpy":
            got = A.to_numpy()
        elif reader == "norm":
            assert relerr(A.MvNorm(tb.NORM_ONE), oracle.norm(want, 0)) <= TOL
            got = A.to_numpy()
        elif reader == "dot":
            B = tb.MultiVec.from_numpy(ctx, a)
            assert relerr(A.MvDot(B), oracle.dot(want, a)) <= TOL
            got = A.to_numpy()
        elif reader == "addmv":
            B = tb.MultiVec(ctx, n, 3)
            B.MvAddMv(1.0, A, 0.0, A)
            got = B.to_numpy()
        elif reader == "apply_view":
            n2, rp, ci, va = tb.galeri("Tridiag", n)
            op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
            Y = tb.MultiVec(ctx, n, 3)
            op.Apply(A, Y)
            assert relerr(Y.to_numpy(), oracle.spmm(rp, ci.astype(np.int32), va, want)) <= TOL
            got = A.to_numpy()
        elif reader == "gram_block":
            assert relerr(A.MvTransMv(1.0, A), oracle.trans_mv(1.0, want, want)) <= TOL
            got = A.to_numpy()
        elif reader == "clonecopy":
            got = A.CloneCopy().to_numpy()
            assert np.array_equal(A.CloneCopy([1]).to_numpy(), want[:, [1]])
        elif reader == "setblock":
            B = tb.MultiVec(ctx, n, 3)
            B.SetBlock(A, [0, 1, 2])
            got = B.to_numpy()
        elif reader == "update_src":
            V = tb.MultiVec.from_numpy(ctx, a[:, [0]])
            bm = rnd(3, 1, 52)
            V.MvTimesMatAddMv(1.0, A, bm, 0.0)
            assert relerr(V.to_numpy(), want @ bm) <= TOL
            got = A.to_numpy()
        elif reader == "scale_again":
            c1.MvScale(4.0)                    # a second factor while the first is pending
            want[:, 1] *= 4.0
            got = A.to_numpy()
        else:
            A.MvInit(3.0)                      # overwriting drops the pending factor
            want[:] = 3.0
            got = A.to_numpy()
        assert np.array_equal(got, want), reader


def test_lazy_copy_survives_every_write_to_its_source(ctx):
    n = 6007
    a, b = rnd(n, 2, 61), rnd(n, 2, 62)
    for write in ("scale", "init", "addmv", "update", "update_beta0", "update_block", "set_numpy", "random", "apply",
                  "setblock", "destroy_source", "none"):
        A = tb.MultiVec.from_numpy(ctx, a)
        B = tb.MultiVec.from_numpy(ctx, b)
        src = A.CloneViewNonConst([1])
        cp = src.CloneCopy()                   # lazy
        if write == "scale":
            src.MvScale(2.0)
        elif write == "init":
            A.MvInit(1.0)
        elif write == "addmv":
            A.MvAddMv(1.0, A, 1.0, B)
        elif write == "update":
            src.MvTimesMatAddMv(-1.0, B, rnd(2, 1, 63), 1.0)      # the save_old path
        elif write == "update_beta0":
            src.MvTimesMatAddMv(-1.0, B, rnd(2, 1, 63), 0.0)
        elif write == "update_block":
            A.MvTimesMatAddMv(-1.0, B, rnd(2, 2, 64), 1.0)
        elif write == "set_numpy":
            A.set_numpy(b)
        elif write == "random":
            A.MvRandom()
        elif write == "apply":
            n2, rp, ci, va = tb.galeri("Tridiag", n)
            op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
            op.Apply(B, A)
        elif write == "setblock":
            A.SetBlock(B, [1, 0])
        elif write == "destroy_source":
            del src
            del A
        assert np.array_equal(cp.to_numpy(), a[:, [1]]), write
        # and a second lazy copy, of the copy
        cp2 = cp.CloneCopy()
        cp.MvScale(0.5)
        assert np.array_equal(cp2.to_numpy(), a[:, [1]]), write
        assert np.array_equal(cp.to_numpy(), a[:, [1]] * 0.5), write


# Fused update + Gram (update_gram.cu): MvTimesMatAddMv also leaves A^T V_new in the mailbox when the call pattern
# predicts that MvTransMv(A, V) comes next (second pass of block Gram-Schmidt, BelosDGKSOrthoManager.hpp:1436-1455).
# Same results as the two separate kernels (<= 1e-13 vs the oracle), never a stale block.
UG_MIN_N = 148 * 28 * 4  # eligibility threshold of update_gram_fits on a 148-SM part


@pytest.mark.parametrize("n", [UG_MIN_N, 20011, 100003, 262147])
@pytest.mark.parametrize("p,q", [(24, 1), (29, 3), (32, 4), (57, 5), (100, 4), (200, 4), (208, 4), (209, 2), (40, 8), (256, 8)])
@pytest.mark.parametrize("alpha,beta", [(-1.0, 1.0), (1.0, 0.0), (0.5, 2.0)])
def test_fused_update_gram(ctx, n, p, q, alpha, beta):
    if (n > 100003 and (p, q) not in [(200, 4), (256, 8), (29, 3)]) or (n == 20011 and alpha != -1.0):
        pytest.skip("covered at the other sizes")
    a, v, bm = rnd(n, p, 21), rnd(n, q, 22), rnd(p, q, 23) * 0.1
    A, V = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, v)
    if beta == 0.0:  # must not read V
        V = tb.MultiVec.from_numpy(ctx, np.full((n, q), np.nan, order="F"))
    # (a one-column CloneCopy is lazy and would be filled by the plain one-column update kernel instead)
    Vold = V.CloneCopy() if q > 1 else None
    ctx.set_fusion(True)  # (resets the call-pattern predictor: the next eligible update speculates)
    l0, g0 = ctx.spec_gram_stats()
    h0 = ctx.dot_hits()
    V.MvTimesMatAddMv(alpha, A, bm, beta)
    G = V.MvTransMv(2.0, A)
    l1, g1 = ctx.spec_gram_stats()
    assert (l1 - l0, g1 - g0) == (1, 1), "the update must have run fused and answered the MvTransMv"
    vnew = oracle.times_mat_add_mv(alpha, a, bm, beta, v)
    assert relerr(V.to_numpy(), vnew) <= TOL
    assert relerr(G, oracle.trans_mv(2.0, a, vnew)) <= TOL
    d_self = V.MvDot(V)
    assert ctx.dot_hits() - h0 == q
    assert relerr(d_self, oracle.dot(vnew, vnew)) <= TOL
    if beta != 0.0 and Vold is not None:
        d_old = V.MvDot(Vold)
        assert ctx.dot_hits() - h0 == 2 * q
        assert np.abs(d_old - oracle.dot(vnew, v)).max() <= TOL * np.sqrt(oracle.dot(vnew, vnew) * oracle.dot(v, v)).max()
    # the separate kernels give the same block
    ctx.set_fusion(False)
    try:
        assert relerr(V.MvTransMv(2.0, A), G) <= TOL
    finally:
        ctx.set_fusion(True)


def test_fused_update_gram_memo_is_never_stale(ctx):
    n, p, q = 50021, 24, 4
    a, v, bm = rnd(n, p, 31), rnd(n, q, 32), rnd(p, q, 33) * 0.1
    A, V = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec.from_numpy(ctx, v)
    ctx.set_fusion(True)
    _, g0 = ctx.spec_gram_stats()
    V.MvTimesMatAddMv(-1.0, A, bm, 1.0)
    vnew = oracle.times_mat_add_mv(-1.0, a, bm, 1.0, v)
    V.MvScale(0.5)  # V changed after the update: the remembered block is for other contents
    assert relerr(V.MvTransMv(1.0, A), oracle.trans_mv(1.0, a, 0.5 * vnew)) <= TOL
    assert ctx.spec_gram_stats()[1] == g0
    # a different view of A (fewer columns) is a different product
    ctx.set_fusion(True)
    V.MvTimesMatAddMv(-1.0, A, bm, 1.0)
    vnew2 = oracle.times_mat_add_mv(-1.0, a, bm, 1.0, 0.5 * vnew)
    Asub = A.CloneView(list(range(p - 1)))
    assert relerr(V.MvTransMv(1.0, Asub), oracle.trans_mv(1.0, a[:, : p - 1], vnew2)) <= TOL
    assert ctx.spec_gram_stats()[1] == g0
    # a column of A rewritten in between
    ctx.set_fusion(True)
    V.MvTimesMatAddMv(-1.0, A, bm, 1.0)
    vnew3 = oracle.times_mat_add_mv(-1.0, a, bm, 1.0, vnew2)
    A.CloneViewNonConst([3]).MvScale(2.0)
    a2 = a.copy()
    a2[:, 3] *= 2.0
    assert relerr(V.MvTransMv(1.0, A), oracle.trans_mv(1.0, a2, vnew3)) <= TOL
    assert ctx.spec_gram_stats()[1] == g0


def test_fused_update_gram_follows_the_two_pass_pattern(ctx):
    """Gram, update, Gram, update per block step (DGKS blkOrtho): after the first steps only the FIRST update of a
    step runs fused (its Gram follows); the second one, whose product nobody asks for, goes to the plain kernel."""
    n, p, q = 60013, 32, 4
    a = rnd(n, p, 41)
    vs = [rnd(n, q, 42), rnd(n, q, 43)]
    A = tb.MultiVec.from_numpy(ctx, a)
    Vs = [tb.MultiVec.from_numpy(ctx, v) for v in vs]
    ctx.set_fusion(True)
    l0, g0 = ctx.spec_gram_stats()
    steps = 12
    for s in range(steps):  # (consecutive block steps work on different blocks)
        V = Vs[s % 2]
        c1 = V.MvTransMv(1.0, A)
        V.MvTimesMatAddMv(-1e-3, A, c1, 1.0)
        c2 = V.MvTransMv(1.0, A)
        V.MvTimesMatAddMv(-1e-3, A, c2, 1.0)
        v = vs[s % 2]
        v = v - 1e-3 * a @ (a.T @ v)
        vs[s % 2] = v - 1e-3 * a @ (a.T @ v)
    l1, g1 = ctx.spec_gram_stats()
    for V, v in zip(Vs, vs):
        assert relerr(V.to_numpy(), v) <= 1e-12
    assert g1 - g0 >= steps - 2, "the second Gram of (almost) every step must come from the fused update"
    assert l1 - l0 <= steps + 3, "the second update of a step must stop speculating after the first steps"


@pytest.mark.parametrize("n", [515, 20011, 100003])
@pytest.mark.parametrize("b,passes", [(2, 1), (4, 1), (4, 2), (8, 1), (8, 2)])
def test_findbasis_after_block_update_needs_no_gram_launch(ctx, n, b, passes):
    """A whole block step as Belos::DGKSOrthoManager / ICGSOrthoManager drive it: the second block update of blkOrtho
    (BelosDGKSOrthoManager.hpp:1454), then findBasis column by column (:1000-1225; ICGS makes two Gram-Schmidt passes
    per column, BelosICGSOrthoManager.hpp:1067).  The cross products of the block update and of the column steps
    (findbasis.cu) answer the one-column MvTransMv calls: no Gram kernel runs (block size 8, whose block update is the
    DMMA kernel without cross products: one, for column 1), values as the oracle's."""
    p = 12
    qb, x, c = rnd(n, p, 61), rnd(n, b, 62), rnd(p, b, 63) * 0.05
    Q, X = tb.MultiVec.from_numpy(ctx, qb), tb.MultiVec.from_numpy(ctx, x)
    ctx.set_fusion(True)
    X.MvTimesMatAddMv(-1.0, Q, c, 1.0)
    ref = oracle.times_mat_add_mv(-1.0, qb, c, 1.0, x)
    h0 = ctx.findbasis_stats()
    ctx.timing_read(reset=True)
    ctx.timing_enable(True)
    for j in range(b):
        Xj = X.CloneViewNonConst([j])
        old = Xj.CloneCopy() if j > 0 else None
        old_dot = Xj.MvDot(Xj)
        assert relerr(old_dot, oracle.dot(ref[:, [j]], ref[:, [j]])) <= TOL
        if j > 0:
            prev = X.CloneView(list(range(j)))
            refold = ref[:, [j]].copy()
            for _ in range(passes):
                prod = Xj.MvTransMv(1.0, prev)
                want = oracle.trans_mv(1.0, ref[:, :j], ref[:, [j]])
                assert np.abs(prod - want).max() <= TOL * np.sqrt(old_dot[0]), (j, prod, want)  # (|prev_i| = 1)
                Xj.MvTimesMatAddMv(-1.0, prev, prod, 1.0)
                ref[:, [j]] = oracle.times_mat_add_mv(-1.0, ref[:, :j], prod, 1.0, ref[:, [j]])
            new_dot = Xj.MvDot(Xj)
            assert relerr(new_dot, oracle.dot(ref[:, [j]], ref[:, [j]])) <= TOL
            od = Xj.MvDot(old)  # (ICGS: the new column with the copy made before both passes -- a kernel of its own there)
            assert np.abs(od - oracle.dot(ref[:, [j]], refold)).max() <= TOL * np.sqrt(new_dot * old_dot).max()
            assert relerr(old.to_numpy(), refold) <= TOL, "the lazy copy must hold the column as it was"
        else:
            new_dot = old_dot
        s = 1.0 / np.sqrt(new_dot[0])
        Xj.MvScale(s)
        ref[:, j] = ref[:, j] * s
    ctx.timing_enable(False)
    t = ctx.timing_read(reset=True)
    grams = (b - 1) * passes
    launched = 1 if b > 4 else 0
    assert ctx.findbasis_stats() - h0 == grams - launched, "the one-column MvTransMv calls must come from the memo"
    assert t["trans_mv"]["calls"] == launched, t
    assert t["dot_norm"]["calls"] == (b - 1 if passes > 1 else 0), t
    assert relerr(X.to_numpy(), ref) <= TOL
    g = oracle.trans_mv(1.0, X.to_numpy(), X.to_numpy())
    assert np.abs(g - np.eye(b)).max() <= 1e-12
