"""SURVEY 8 f.3 -- the TSQR hooks of Belos::MultiVec (factorExplicit / revealRank, belos/src/BelosMultiVec.hpp:213-293)
on the GPU, and Belos::TsqrOrthoManager (unmodified, belos/src/BelosTsqrOrthoManagerImpl.hpp) driven through them.

Oracle: Householder QR of LAPACK (numpy.linalg.qr) with R's diagonal made positive -- for a full-rank block the
factorisation A = QR with positive diagonal is unique, so R must agree entry for entry up to rounding.  The
reference's own TSQR (Kokkos) is not buildable here: parity for this row is unpinned by reference output."""
import numpy as np
import pytest

import trilinos_b200 as tb

pytestmark = pytest.mark.gpu
EPS = np.finfo(float).eps


def rnd(n, k, seed):
    return np.asfortranarray(np.random.default_rng(seed).uniform(-1, 1, (n, k)))


def qr_pos(a):
    q, r = np.linalg.qr(a)
    s = np.sign(np.diag(r))
    s[s == 0] = 1.0
    return q * s, (r.T * s).T


@pytest.mark.parametrize("n,b", [(1000, 1), (5003, 4), (20000, 8), (100003, 16), (3001, 24)])
def test_factor_explicit_matches_householder_qr(ctx, n, b):
    a = rnd(n, b, 7 + b)
    A, Q = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec(ctx, n, b)
    R = A.factorExplicit(Q)
    q = Q.to_numpy()
    qh, rh = qr_pos(a)
    assert np.array_equal(A.to_numpy(), a), "A must not be modified"
    assert np.allclose(np.tril(R, -1), 0.0, atol=0.0) and np.all(np.diag(R) > 0)
    assert np.abs(q.T @ q - np.eye(b)).max() <= 50 * EPS * np.sqrt(n)
    assert np.linalg.norm(a - q @ R) <= 1e-13 * np.linalg.norm(a)
    assert np.linalg.norm(R - rh) <= 1e-12 * np.linalg.norm(rh)
    assert np.linalg.norm(q - qh) <= 1e-11 * np.sqrt(b)


def test_factor_explicit_ill_conditioned_block(ctx):
    """cond(A) ~ 1e10: plain Cholesky QR breaks down, the shifted three-pass variant must not."""
    n, b = 20011, 6
    u, _ = np.linalg.qr(rnd(n, b, 3))
    v, _ = np.linalg.qr(rnd(b, b, 4))
    a = np.asfortranarray((u * np.logspace(0, -10, b)) @ v.T)
    A, Q = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec(ctx, n, b)
    R = A.factorExplicit(Q)
    q = Q.to_numpy()
    assert np.abs(q.T @ q - np.eye(b)).max() <= 1e-11
    assert np.linalg.norm(a - q @ R) <= 1e-12 * np.linalg.norm(a)
    assert Q.revealRank(np.asfortranarray(R.copy()), 1e-13) == b


def test_reveal_rank_on_deficient_block(ctx):
    n, b = 9001, 6
    a = rnd(n, b, 5)
    a[:, 2] = a[:, 0] - 2.0 * a[:, 1]     # dependent column
    a[:, 5] = 0.0                          # zero column
    A, Q = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec(ctx, n, b)
    R = np.asfortranarray(A.factorExplicit(Q))
    q0 = Q.to_numpy()
    assert np.abs(q0.T @ q0 - np.eye(b)).max() <= 1e-10, "Q must be orthonormal even for a rank-deficient block"
    assert np.linalg.norm(a - q0 @ R) <= 1e-11 * np.linalg.norm(a)
    rank = Q.revealRank(R, 1e-10)
    assert rank == 4
    q = Q.to_numpy()
    assert np.abs(q.T @ q - np.eye(b)).max() <= 1e-10
    assert np.linalg.norm(a - q @ R) <= 1e-11 * np.linalg.norm(a)          # A = (Q U)(S V^T)
    # the first `rank` columns of Q span range(A)
    proj = q[:, :rank] @ (q[:, :rank].T @ a)
    assert np.linalg.norm(a - proj) <= 1e-10 * np.linalg.norm(a)
    assert np.abs(R[rank:, :]).max() <= 1e-9 * np.abs(R).max()


def test_reveal_rank_full_rank_leaves_q_and_r_alone(ctx):
    n, b = 4000, 5
    a = rnd(n, b, 9)
    A, Q = tb.MultiVec.from_numpy(ctx, a), tb.MultiVec(ctx, n, b)
    R = np.asfortranarray(A.factorExplicit(Q))
    q0, r0 = Q.to_numpy(), R.copy()
    assert Q.revealRank(R, 1e-12) == b
    assert np.array_equal(Q.to_numpy(), q0) and np.array_equal(R, r0)


@pytest.mark.parametrize("ortho", ["TSQR", "DGKS", "ICGS", "IMGS"])
@pytest.mark.parametrize("n,ncols,x1,x2", [(500, 12, 3, 5), (20000, 8, 4, 4)])
def test_ortho_manager_bounds(ctx, ortho, n, ncols, x1, x2):
    """Belos::TsqrOrthoManager / DGKS / ICGS / IMGS (the reference's classes, made by Belos::OrthoManagerFactory) on
    B200MultiVec blocks: the normalize / projectAndNormalize checks of belos/src/BelosOrthoManagerTest.hpp on full-rank
    blocks.  Rank-deficient blocks are NOT driven through TsqrOrthoManager: its null-space replacement takes
    CloneViewNonConst(Q, Range1D(rank, ncols-1)) (BelosTsqrOrthoManagerImpl.hpp:1770-1777), and the reference's
    MultiVecTraits<S, Belos::MultiVec<S>> overload for Range1D ignores the lower bound (BelosMultiVec.hpp:388-398:
    indVec[k] = k) -- with the run-time polymorphic interface the reference itself views the wrong columns there
    (here: caught as aliasing by MvTimesMatAddMv).  The hooks' own rank revelation is tested above."""
    assert tb.ortho_test(ctx, ortho, n=n, ncols=ncols, size_x1=x1, size_x2=x2, rank_revealing=False) == 0


def test_pseudo_block_gmres_with_tsqr(ctx):
    """"Orthogonalization": "TSQR" is accepted by PseudoBlockGmresSolMgr under HAVE_BELOS_TSQR
    (belos/src/BelosPseudoBlockGmresSolMgr.hpp:686-731); same iteration count as DGKS on the same problem."""
    import oracle
    n, rp, ci, va = tb.galeri("Laplace3D", 12, 12, 12)
    op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
    xe, _ = oracle.random(1, n, 3)
    Xe = tb.MultiVec.from_numpy(ctx, xe)
    B = tb.MultiVec(ctx, n, 3)
    op.Apply(Xe, B)
    its = {}
    for ortho in ("DGKS", "TSQR"):
        X = tb.MultiVec(ctx, n, 3)
        r = tb.PseudoBlockGmresSolMgr(op, block_size=3, num_blocks=40, max_iters=-1, max_restarts=50, tol=1e-8, ortho=ortho).solve(X, B)
        assert r["converged"] and max(r["true_rel_resid"]) <= 1e-8 * 1.001, (ortho, r)
        its[ortho] = r["num_iters"]
    assert abs(its["TSQR"] - its["DGKS"]) <= 1, its
