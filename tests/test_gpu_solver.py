"""GPU tests of the whole path: the reference's own contract test (Belos::TestMultiVecTraits /
TestOperatorTraits) on the B200 plugin classes, and the unmodified Belos::BlockGmresSolMgr solving
the golden problems — iteration counts within +-2 % of the reference Epetra/Belos run on the same
explicit right-hand sides (tests/golden/ref_solves*.json, produced by oracle/gen_golden.py).
"""
import json
import os

import numpy as np
import pytest

import oracle
import trilinos_b200 as tb

pytestmark = pytest.mark.gpu
GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")
SOLVES = json.load(open(os.path.join(GOLD, "ref_solves.json")))
FULL = json.load(open(os.path.join(GOLD, "ref_solves_full.json")))
MANAGERS = json.load(open(os.path.join(GOLD, "ref_solves_managers.json")))


@pytest.mark.parametrize("n,k", [(100, 5), (1000, 10), (37, 3)])
def test_belos_mvop_tester(ctx, n, k):
    """belos/epetra/test/MVOPTester/cxx_main.cpp:91-176 with B200MultiVec / B200CrsOp (dim 100, 5 columns, tridiagonal A)."""
    assert tb.mvop_test(ctx, n, k)


def run_case(ctx, g, history=0):
    name, dims, b = g["matrix"], g["dims"], g["block_size"]
    n, rp, ci, va = tb.galeri(name, *dims)
    op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
    xe, _ = oracle.random(1, n, b)          # bit-identical to the reference's Epetra LCG stream
    Xe = tb.MultiVec.from_numpy(ctx, xe)
    B = tb.MultiVec(ctx, n, b)
    op.Apply(Xe, B)
    X = tb.MultiVec(ctx, n, b)
    mgr = tb.SOLVER_MANAGERS[g.get("solver", "BlockGmres")](op, block_size=b, num_blocks=g["num_blocks"], max_iters=g.get("max_iters", -1),
                              max_restarts=g["max_restarts"], tol=g["tol"], ortho=g["ortho"])
    r = mgr.solve(X, B, history=history) if history else mgr.solve(X, B)
    return r, X.to_numpy(), xe


# Relative tolerance on the native residual norms of a block iteration, GPU run vs reference run of the same solve.
# The two runs differ only in summation order (1e-16-level perturbations of every inner product); over a restart
# cycle GMRES amplifies those by at most the conditioning of the least-squares problem, far below this bound while
# the residual has not stagnated.  Measured: see profiles/r2_fullsize_parity.txt.
HIST_RTOL = 1e-8


def check_history(r, g, rows=None):
    ref = np.array(g["native_resid_history"])
    got = np.array(r["native_resid_history"])
    m = min(len(ref), len(got)) if rows is None else rows
    assert m >= min(len(ref), 33), (len(ref), len(got))
    rel = np.abs(got[:m] - ref[:m]) / np.maximum(np.abs(ref[:m]), 1e-300)
    assert rel.max() <= HIST_RTOL, (float(rel.max()), np.unravel_index(rel.argmax(), rel.shape))
    return float(rel.max())


@pytest.mark.parametrize("label", sorted(SOLVES))
def test_iteration_count_parity_small(ctx, label):
    g = SOLVES[label]
    r, x, xe = run_case(ctx, g)
    assert r["converged"] == g["converged"]
    assert max(r["true_rel_resid"]) <= g["tol"] * 1.001
    assert abs(r["num_iters"] - g["num_iters"]) <= max(1, round(0.02 * g["num_iters"])), (r["num_iters"], g["num_iters"])
    assert np.abs(x - xe).max() <= 10 * max(g["max_abs_err"], 1e-12)


@pytest.mark.parametrize("label", sorted(FULL))
def test_iteration_count_parity_baseline_configs(ctx, label):
    """All five BASELINE configs at FULL size on one GPU against runs of the reference's own Epetra/Belos code
    (oracle/gen_golden.py --full / --configs345): configs[0] Laplace2D 100^2 b=1, configs[1] Laplace3D 100^3 b=4,
    configs[2] Laplace3D 200^3 b=4 (8M rows), configs[3] Brick3D 150^3 b=8 ICGS -- iteration count within +-2 %,
    same achieved tolerance, and (where the golden run recorded it) the native residual norms of the first restart
    cycle within HIST_RTOL; configs[4] Recirc2D 2000^2 b=16 m=100 -- a bounded run on both sides: the first restart
    cycle on the GPU (100 block iterations) against the reference's first 32 (single-rank Epetra cannot allocate more
    than 528 basis columns of 4M rows: `new double[int * int]`, Epetra_MultiVector.cpp:257; block Arnoldi does not depend
    on the restart length before the first restart, so its m = 32 run is the start of the same cycle): residual history
    within HIST_RTOL for every recorded iteration."""
    g = FULL[label]
    bounded = not g["converged"]                 # config 5: the reference run is its first restart cycle only
    hist = g.get("native_resid_history")
    r, x, xe = run_case(ctx, g, history=(len(hist) + 8) if hist else 0)
    if bounded:
        assert not r["converged"] and r["num_iters"] == g["max_iters"], (r["num_iters"], g["max_iters"])
        assert len(hist) == g["num_iters"] + 1
        check_history(r, g, rows=len(hist))
        return
    assert r["converged"]
    assert max(r["true_rel_resid"]) <= g["tol"] * 1.001
    assert abs(r["num_iters"] - g["num_iters"]) <= max(1, round(0.02 * g["num_iters"])), (r["num_iters"], g["num_iters"])
    assert abs(r["achieved_tol"] - g["achieved_tol"]) <= 1e-3 * g["achieved_tol"] + 1e-12
    if hist:
        check_history(r, g, rows=min(len(hist), g["num_blocks"] + 1))


@pytest.mark.parametrize("label", sorted(MANAGERS))
def test_other_solver_managers_iteration_parity(ctx, label):
    """Belos::PseudoBlockGmresSolMgr / BlockCGSolMgr / PseudoBlockCGSolMgr (unmodified, instantiated on the
    plugin classes) against the reference's Epetra runs of the same managers; same +-2 % iteration criterion."""
    g = MANAGERS[label]
    r, x, xe = run_case(ctx, g)
    assert r["converged"]
    assert max(r["true_rel_resid"]) <= g["tol"] * 1.001
    assert abs(r["num_iters"] - g["num_iters"]) <= max(1, round(0.02 * g["num_iters"])), (r["num_iters"], g["num_iters"])
    assert np.abs(x - xe).max() <= 10 * max(g["max_abs_err"], 1e-12)


def test_known_answer_diagonal(ctx):
    """belos/epetra/test/BlockGmres/test_bl_gmres_diag.cpp:355-411 in spirit: diagonal operator D = diag(1..n),
    B = ones -> X = 1/i, ||X - X_exact|| < 1e-10."""
    n, b = 10, 2
    n_, rp, ci, va = tb.galeri("Diag", n, conv=0.0)
    op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
    B = tb.MultiVec(ctx, n, b)
    B.MvInit(1.0)
    X = tb.MultiVec(ctx, n, b)
    mgr = tb.BlockGmresSolMgr(op, block_size=1, num_blocks=n, max_iters=50, max_restarts=5, tol=1e-12)
    r = mgr.solve(X, B)
    assert r["converged"]
    x = X.to_numpy()
    assert np.linalg.norm(x[:, 0] - 1.0 / np.arange(1, n + 1)) < 1e-10
    assert np.linalg.norm(x[:, 1] - 1.0 / np.arange(1, n + 1)) < 1e-10


def test_larger_block_than_rhs_and_ortho_variants(ctx):
    """--num-rhs=1 --block-size=2 (augmented block, belos/epetra/test/BlockGmres/CMakeLists.txt) and IMGS."""
    n, rp, ci, va = tb.galeri("Laplace3D", 12, 12, 12)
    op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
    xe, _ = oracle.random(3, n, 1)
    Xe = tb.MultiVec.from_numpy(ctx, xe)
    B = tb.MultiVec(ctx, n, 1)
    op.Apply(Xe, B)
    for ortho, bs in (("DGKS", 2), ("IMGS", 1), ("ICGS", 2)):
        X = tb.MultiVec(ctx, n, 1)
        r = tb.BlockGmresSolMgr(op, block_size=bs, num_blocks=40, max_iters=-1, max_restarts=50, tol=1e-9, ortho=ortho).solve(X, B)
        assert r["converged"], (ortho, bs, r)
        assert max(r["true_rel_resid"]) <= 1e-9 * 1.001
