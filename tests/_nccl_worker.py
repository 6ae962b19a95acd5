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
    # Irregular neighbour relations (every rank must walk through the same collective steps whatever its own rows couple
    # to): (i) an upper-bidiagonal operator -- rank r needs one row of rank r+1 and nothing of rank r-1, so the halo
    # relation is not mutual and the last rank has no ghosts at all; (ii) a tridiagonal operator whose first rank block is
    # cut off from the rest -- rank 0 has no halo neighbour while (world >= 3) the others exchange.
    from trilinos_b200.core import linear_partition
    for case in ("bidiagonal", "rank0_decoupled"):
        n = 41 * world + 5
        lo, hi = linear_partition(n, world, rank)
        cut = linear_partition(n, world, 0)[1]  # first row of rank 1
        rows = []
        for i in range(n):
            ent = [(i, 4.0 + 0.01 * i)]
            if i + 1 < n and not (case == "rank0_decoupled" and i + 1 == cut):
                ent.append((i + 1, -1.0 - 0.001 * i))
            if case == "rank0_decoupled" and i > 0 and i != cut:
                ent.insert(0, (i - 1, -0.5))
            rows.append(ent)
        rp = np.zeros(n + 1, dtype=np.int32)
        for i, ent in enumerate(rows):
            rp[i + 1] = rp[i] + len(ent)
        ci = np.array([c for ent in rows for c, _ in ent], dtype=np.int32)
        va = np.array([v for ent in rows for _, v in ent], dtype=np.float64)
        lrp = (rp[lo:hi + 1] - rp[lo]).astype(np.int32)
        op = tb.CrsOperator(ctx, n, lo, hi, lrp, ci[rp[lo]:rp[hi]], va[rp[lo]:rp[hi]])
        k = 3
        xg = np.asfortranarray(np.random.default_rng(11).uniform(-1, 1, (n, k)))
        X = tb.MultiVec.from_numpy(ctx, xg[lo:hi], n_global=n, row_offset=lo)
        Y = tb.dist.local_multivec(ctx, n, lo, hi, k)
        yg = oracle.spmm(rp, ci, va, xg)
        for _ in range(3):  # (back-to-back applies: window parity, exchange numbering)
            op.Apply(X, Y)
        errs = dict(apply=rel(Y.to_numpy(), yg[lo:hi]), dot=rel(Y.MvDot(X), oracle.dot(yg, xg)),
                    n2=rel(Y.MvNorm(tb.NORM_TWO), oracle.norm(yg, 1)), gram=rel(Y.MvTransMv(1.0, X), oracle.trans_mv(1.0, xg, yg)))
        assert max(errs.values()) <= TOL, (case, errs)
        out[case] = errs
    # one block step of DGKS (BelosDGKSOrthoManager.hpp:1400-1472, :1000-1225) on row-partitioned random data, at a size where
    # the fused update + Gram kernel and the findBasis memo are active: Gram, update (+ the next Gram from the mailbox),
    # update (+ cross products), then the column steps -- every value against the serial oracle
    n, p, b = 18000 * world + 7, 40, 4
    lo, hi = linear_partition(n, world, rank)
    rng = np.random.default_rng(7)
    qg = np.asfortranarray(rng.uniform(-1, 1, (n, p)))
    Q = tb.MultiVec.from_numpy(ctx, qg[lo:hi], n_global=n, row_offset=lo)
    ctx.set_fusion(True)
    for rnd_ in range(2):  # (second round: the call-pattern predictor has learned that only the first update is followed by a Gram)
        xg = np.asfortranarray(rng.uniform(-1, 1, (n, b)))
        X = tb.MultiVec.from_numpy(ctx, xg[lo:hi], n_global=n, row_offset=lo)
        s0, g0 = ctx.spec_gram_stats()
        f0 = ctx.findbasis_stats()
        errs = {}
        c1 = X.MvTransMv(1.0, Q)
        errs["gram1"] = rel(c1, oracle.trans_mv(1.0, qg, xg))
        c1 = c1 * 1e-3
        X.MvTimesMatAddMv(-1.0, Q, c1, 1.0)
        xg = oracle.times_mat_add_mv(-1.0, qg, c1, 1.0, xg)
        c2 = X.MvTransMv(1.0, Q)
        errs["gram2_from_fused_update"] = rel(c2, oracle.trans_mv(1.0, qg, xg))
        c2 = c2 * 1e-3
        X.MvTimesMatAddMv(-1.0, Q, c2, 1.0)
        xg = oracle.times_mat_add_mv(-1.0, qg, c2, 1.0, xg)
        s1, g1 = ctx.spec_gram_stats()
        assert g1 - g0 == 1, "the second Gram must have come from the fused update"
        for j in range(b):
            Xj = X.CloneViewNonConst([j])
            old = Xj.CloneCopy() if j > 0 else None
            d0 = Xj.MvDot(Xj)
            errs[f"olddot{j}"] = rel(d0, oracle.dot(xg[:, [j]], xg[:, [j]]))
            if j > 0:
                prev = X.CloneView(list(range(j)))
                prod = Xj.MvTransMv(1.0, prev)
                want = oracle.trans_mv(1.0, xg[:, :j], xg[:, [j]])
                errs[f"prod{j}"] = float(np.abs(prod - want).max() / np.sqrt(d0[0]))
                Xj.MvTimesMatAddMv(-1.0, prev, prod, 1.0)
                xg[:, [j]] = oracle.times_mat_add_mv(-1.0, xg[:, :j], prod, 1.0, xg[:, [j]])
                d1 = Xj.MvDot(Xj)
                errs[f"newdot{j}"] = rel(d1, oracle.dot(xg[:, [j]], xg[:, [j]]))
                old.to_numpy()  # (a lazy copy that is read: filled by the column step)
            else:
                d1 = d0
            sc = 1.0 / np.sqrt(d1[0])
            Xj.MvScale(sc)
            xg[:, j] *= sc
        errs["x"] = rel(X.to_numpy(), xg[lo:hi])
        assert max(errs.values()) <= TOL, errs
        fb = ctx.findbasis_stats() - f0
        if rnd_ == 1:
            assert s1 - s0 == 1 and fb == b - 1, (s1 - s0, fb)
        out[f"block_step_{rnd_}"] = dict(max_err=max(errs.values()), spec_launches=s1 - s0, gram_memo_hits=g1 - g0,
                                         findbasis_hits=fb)
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
