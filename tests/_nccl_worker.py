"""Worker for tests/test_multigpu_nccl.py (launched by torch.distributed.run, one rank per GPU).

Every rank owns its row block of the linear map (Epetra_Map(n,0,comm), epetra/src/Epetra_BlockMap.cpp:83-90),
runs the product's distributed path (NCCL halo exchange + allreduced reductions + the unmodified
Belos::BlockGmresSolMgr) and compares with the SERIAL oracle on the same global inputs — the same
check the reference's MPI tests make implicitly (partition-independent results up to summation order).
"""
import json
import os
import sys

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import oracle  # noqa: E402
import trilinos_b200 as tb  # noqa: E402

TOL = 1e-13


def rel(a, r):
    return float(np.linalg.norm(np.ravel(a - r)) / max(np.linalg.norm(np.ravel(r)), 1e-300))


def main():
    ctx, rank, world = tb.dist.init_from_env()
    gold = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_solves.json")))
    out = {}
    for name, dims, k in (("Laplace3D", (24, 24, 24), 4), ("Recirc2D", (41, 37, 1), 3), ("Brick3D", (9, 8, 11), 8),
                          ("Laplace2D", (100, 100, 1), 1)):
        op, n, lo, hi = tb.dist.local_operator(ctx, name, dims)
        ng, rp, ci, va = oracle.galeri_csr(name, *dims)
        assert ng == n
        xg, _ = oracle.random(1, n, k)                 # global multivector, Epetra LCG stream
        ctx.set_seed(1)
        X = tb.dist.local_multivec(ctx, n, lo, hi, k, zero=False)
        X.MvRandom()                                   # partition-independent: must equal rows [lo,hi) of xg
        assert np.array_equal(X.to_numpy(), xg[lo:hi]), f"{name}: MvRandom is not partition independent"
        Y = tb.dist.local_multivec(ctx, n, lo, hi, k)
        yg = oracle.spmm(rp, ci.astype(np.int32), va, xg)
        # one column first: the halo window is then re-made (collectively) for the wider block below
        op.Apply(X.CloneView([k - 1]), Y.CloneViewNonConst([0]))
        assert rel(Y.to_numpy()[:, 0], yg[lo:hi, k - 1]) <= TOL, f"{name}: single-column apply"
        op.Apply(X, Y)
        op.Apply(X, Y)                                  # back-to-back applies: ghost-block parity alternates
        e_apply = rel(Y.to_numpy(), yg[lo:hi])
        e_gram = rel(Y.MvTransMv(1.0, X), oracle.trans_mv(1.0, xg, yg))
        e_dot = rel(Y.MvDot(X), oracle.dot(yg, xg))
        e_n2 = rel(Y.MvNorm(tb.NORM_TWO), oracle.norm(yg, 1))
        e_n1 = rel(Y.MvNorm(tb.NORM_ONE), oracle.norm(yg, 0))
        e_ni = rel(Y.MvNorm(tb.NORM_INF), oracle.norm(yg, 2))
        errs = dict(apply=e_apply, gram=e_gram, dot=e_dot, n2=e_n2, n1=e_n1, ninf=e_ni)
        assert max(errs.values()) <= TOL, (name, errs)
        out[name] = errs
    # solves: iteration counts against the reference's golden runs (single-rank Epetra/Belos)
    for label, g in sorted(gold.items()):
        name, dims, b = g["matrix"], tuple(g["dims"]), g["block_size"]
        op, n, lo, hi = tb.dist.local_operator(ctx, name, dims)
        ctx.set_seed(1)
        Xe = tb.dist.local_multivec(ctx, n, lo, hi, b, zero=False)
        Xe.MvRandom()
        B = tb.dist.local_multivec(ctx, n, lo, hi, b)
        op.Apply(Xe, B)
        X = tb.dist.local_multivec(ctx, n, lo, hi, b)
        mgr = tb.BlockGmresSolMgr(op, block_size=b, num_blocks=g["num_blocks"], max_iters=g.get("max_iters", -1),
                                  max_restarts=g["max_restarts"], tol=g["tol"], ortho=g["ortho"])
        r = mgr.solve(X, B)
        assert r["converged"] == g["converged"], (label, r)
        assert max(r["true_rel_resid"]) <= g["tol"] * 1.001, (label, r)
        assert abs(r["num_iters"] - g["num_iters"]) <= max(1, round(0.02 * g["num_iters"])), (label, r["num_iters"], g["num_iters"])
        assert np.abs(X.to_numpy() - Xe.to_numpy()).max() <= 10 * max(g["max_abs_err"], 1e-12)
        out[label] = dict(iters=r["num_iters"], golden=g["num_iters"])
    # other solver managers and the Chebyshev right preconditioner on the distributed path
    more = json.load(open(os.path.join(ROOT, "tests", "golden", "ref_solves_managers.json")))
    more.update(json.load(open(os.path.join(ROOT, "tests", "golden", "ref_solves_prec.json"))))
    for label, g in sorted(more.items()):
        name, dims, b = g["matrix"], tuple(g["dims"]), g["block_size"]
        op, n, lo, hi = tb.dist.local_operator(ctx, name, dims)
        ctx.set_seed(1)
        Xe = tb.dist.local_multivec(ctx, n, lo, hi, b, zero=False)
        Xe.MvRandom()
        B = tb.dist.local_multivec(ctx, n, lo, hi, b)
        op.Apply(Xe, B)
        X = tb.dist.local_multivec(ctx, n, lo, hi, b)
        ch = g.get("cheby")
        P = tb.Chebyshev(op, degree=ch["degree"], lambda_max=ch.get("lambda_max", -1.0), seed=1) if ch else None
        mgr = tb.SOLVER_MANAGERS[g["solver"]](op, block_size=b, num_blocks=g["num_blocks"], max_iters=g.get("max_iters", -1),
                                              max_restarts=g["max_restarts"], tol=g["tol"], ortho=g["ortho"], right_prec=P)
        r = mgr.solve(X, B)
        assert r["converged"], (label, r)
        assert max(r["true_rel_resid"]) <= g["tol"] * 1.001, (label, r)
        assert abs(r["num_iters"] - g["num_iters"]) <= max(1, round(0.02 * g["num_iters"])), (label, r["num_iters"], g["num_iters"])
        out[label] = dict(iters=r["num_iters"], golden=g["num_iters"])
    ctx.sync()
    if rank == 0:
        print("NCCL_WORKER_OK " + json.dumps(dict(world=world, results=out)))
    import torch.distributed as dist
    if dist.is_initialized():
        dist.barrier()
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
