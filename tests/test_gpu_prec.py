"""GPU tests of SURVEY 8 f.1: right preconditioning through the same boundary.  The Chebyshev smoother
(b200_cheby_*, Belos::B200ChebyshevOp) against the reference's Ifpack_Chebyshev (fixtures: oracle/gen_golden.py --cheby)
and the unmodified Belos solver managers with it as right preconditioner against the reference's
Belos::EpetraPrecOp(Ifpack_Chebyshev) runs.  Tolerances: 1e-13 normwise per apply, +-2 % iterations."""
import json
import os

import numpy as np
import pytest

import oracle
import trilinos_b200 as tb
from oracle.gen_golden import CHEBY_APPLY

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
K = np.load(os.path.join(GOLD, "ref_cheby.npz"))
PREC = json.load(open(os.path.join(GOLD, "ref_solves_prec.json")))


def rel(a, r):
    return np.linalg.norm(np.ravel(a - r)) / max(np.linalg.norm(np.ravel(r)), 1e-300)


def make_op(ctx, name, dims):
    kw = dict(conv=1.0, diff=1e-5) if name == "Recirc2D" else {}
    n, rp, ci, va = tb.galeri(name, *dims, **kw)
    return n, tb.CrsOperator(ctx, n, 0, n, rp, ci, va), (rp, ci, va)


@pytest.mark.parametrize("case", CHEBY_APPLY, ids=[c[0] for c in CHEBY_APPLY])
def test_chebyshev_apply_parity(ctx, case):
    label, name, dims, k, degree, lam, ratio = case
    n, op, _ = make_op(ctx, name, dims)
    X = tb.MultiVec.from_numpy(ctx, K[label + "_X"])
    Y = tb.MultiVec(ctx, n, k)
    # (a) the reference's eigenvalue handed over: same coefficients and operation order in the smoother passes
    #     (bit-identical for degree 1; from degree 2 on the operator apply in between uses FMA, the reference does not)
    P = tb.Chebyshev(op, degree=degree, lambda_max=float(K[label + "_lambda"]), eig_ratio=ratio)
    P.ApplyInverse(X, Y)
    if degree == 1:
        assert np.array_equal(Y.to_numpy(), K[label + "_Y"])
    assert rel(Y.to_numpy(), K[label + "_Y"]) <= 1e-13
    # (b) own power method (Ifpack_Chebyshev::PowerMethod, start vector = Epetra LCG seed 1)
    if lam < 0:
        P2 = tb.Chebyshev(op, degree=degree, lambda_max=-1.0, eig_ratio=ratio, seed=1)
        lmax, lmin = P2.lambdas()
        assert abs(lmax - float(K[label + "_lambda"])) <= 1e-12 * lmax
        assert abs(lmin - lmax / ratio) <= 1e-15
        P2.ApplyInverse(X, Y)
        assert rel(Y.to_numpy(), K[label + "_Y"]) <= 1e-12


def test_chebyshev_aliased_and_views(ctx):
    """X and Y the same memory (the AztecOO case, Ifpack_Chebyshev.cpp:482-488) and non-contiguous column views."""
    n, op, (rp, ci, va) = make_op(ctx, "Laplace3D", (9, 8, 7))
    x, _ = oracle.random(3, n, 5)
    ref = oracle.cheby_apply(rp, ci.astype(np.int32), va, x[:, [4, 1, 2]], 4, 1.8)
    P = tb.Chebyshev(op, degree=4, lambda_max=1.8)
    X = tb.MultiVec.from_numpy(ctx, x)
    Y = tb.MultiVec(ctx, n, 5)
    P.ApplyInverse(X.CloneView([4, 1, 2]), Y.CloneViewNonConst([0, 2, 3]))
    assert rel(Y.to_numpy()[:, [0, 2, 3]], ref) <= 1e-13
    assert not Y.to_numpy()[:, [1, 4]].any()
    Xv = X.CloneViewNonConst([4, 1, 2])
    P.ApplyInverse(Xv, Xv)
    assert rel(X.to_numpy()[:, [4, 1, 2]], ref) <= 1e-13
    assert np.array_equal(X.to_numpy()[:, [0, 3]], x[:, [0, 3]])
    with pytest.raises(tb.B200Error):
        tb.Chebyshev(op, degree=0)


def test_elementwise_multiply_and_diagonal(ctx):
    """Epetra_MultiVector::Multiply(ScalarAB, A, B, ScalarThis) (element-wise) and Epetra_CrsMatrix::ExtractDiagonalCopy."""
    n, op, (rp, ci, va) = make_op(ctx, "Recirc2D", (12, 9, 1))
    d = op.ExtractDiagonalCopy().to_numpy()[:, 0]
    rows = np.repeat(np.arange(n), np.diff(rp))
    dref = np.zeros(n)
    dref[rows[ci == rows]] = va[ci == rows]
    assert np.array_equal(d, dref)
    b, _ = oracle.random(5, n, 3)
    y0, _ = oracle.random(6, n, 3)
    D = tb.MultiVec.from_numpy(ctx, d[:, None])
    B = tb.MultiVec.from_numpy(ctx, b)
    for sab, sthis in ((1.0, 0.0), (-0.5, 0.0), (2.0, 1.0), (0.25, -3.0)):
        Y = tb.MultiVec.from_numpy(ctx, y0)
        Y.Multiply(sab, D, B, sthis)
        assert np.array_equal(Y.to_numpy(), sthis * y0 + (sab * d)[:, None] * b)
    Y = tb.MultiVec.from_numpy(ctx, np.full((n, 3), np.nan))
    Y.Multiply(1.0, D, B, 0.0)          # ScalarThis == 0 never reads this
    assert np.isfinite(Y.to_numpy()).all()


@pytest.mark.parametrize("label", sorted(PREC))
def test_right_preconditioned_solves(ctx, label):
    g = PREC[label]
    n, op, _ = make_op(ctx, g["matrix"], g["dims"])
    b = g["block_size"]
    xe, _ = oracle.random(1, n, b)
    Xe = tb.MultiVec.from_numpy(ctx, xe)
    B = tb.MultiVec(ctx, n, b)
    op.Apply(Xe, B)
    X = tb.MultiVec(ctx, n, b)
    ch = g["cheby"]
    P = tb.Chebyshev(op, degree=ch["degree"], lambda_max=ch.get("lambda_max", -1.0), seed=1)
    mgr = tb.SOLVER_MANAGERS[g["solver"]](op, block_size=b, num_blocks=g["num_blocks"], max_iters=-1,
                                          max_restarts=g["max_restarts"], tol=g["tol"], ortho=g["ortho"], right_prec=P)
    r = mgr.solve(X, B)
    assert r["converged"]
    assert max(r["true_rel_resid"]) <= g["tol"] * 1.001
    assert abs(r["num_iters"] - g["num_iters"]) <= max(1, round(0.02 * g["num_iters"])), (r["num_iters"], g["num_iters"])
    assert np.abs(X.to_numpy() - xe).max() <= 10 * max(g["max_abs_err"], 1e-12)


def test_full_size_preconditioned_config2(ctx):
    """BASELINE configs[1] (Laplace3D 100^3, b = 4, 50 blocks, DGKS) with a degree-8 Chebyshev right preconditioner:
    size-independent properties (converged, true residual <= tol, far fewer iterations than the 289 unpreconditioned)."""
    n, op, _ = make_op(ctx, "Laplace3D", (100, 100, 100))
    ctx.set_seed(1)
    Xe = tb.MultiVec(ctx, n, 4, zero=False)
    Xe.MvRandom()
    B = tb.MultiVec(ctx, n, 4)
    op.Apply(Xe, B)
    X = tb.MultiVec(ctx, n, 4)
    P = tb.Chebyshev(op, degree=8)
    r = tb.BlockGmresSolMgr(op, block_size=4, num_blocks=50, max_iters=-1, max_restarts=1000, tol=1e-8, ortho="DGKS",
                            right_prec=P).solve(X, B)
    assert r["converged"] and max(r["true_rel_resid"]) <= 1e-8 * 1.001
    assert r["num_iters"] < 100
