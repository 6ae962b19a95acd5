"""Small pass through the round-2 kernels (fused update + Gram in both update roles, findBasis column steps, window
SpMM) against the oracle -- the case meant for compute-sanitizer (closed on this GPU pool: the plain run is kept as a
quick check).  usage: python tools/sanitize_case.py"""
import sys
import numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb
import oracle

ctx = tb.Context(0)
rng = np.random.default_rng(5)
def rel(a, r):
    return float(np.linalg.norm(np.ravel(a - r)) / max(np.linalg.norm(np.ravel(r)), 1e-300))
for n, p, q in ((16600, 40, 4), (16700, 100, 4), (16650, 50, 8)):
    qa = np.asfortranarray(rng.uniform(-1, 1, (n, p))); xa = np.asfortranarray(rng.uniform(-1, 1, (n, q)))
    Q, X = tb.MultiVec.from_numpy(ctx, qa), tb.MultiVec.from_numpy(ctx, xa)
    ctx.set_fusion(True)
    c1 = X.MvTransMv(1.0, Q) * 1e-3
    X.MvTimesMatAddMv(-1.0, Q, c1, 1.0)
    xn = oracle.times_mat_add_mv(-1.0, qa, c1, 1.0, xa)
    c2 = X.MvTransMv(1.0, Q)
    print(n, p, q, "fused:", ctx.spec_gram_stats(), rel(X.to_numpy(), xn), rel(c2, oracle.trans_mv(1.0, qa, xn)))
    for j in range(1, min(q, 4)):
        Xj = X.CloneViewNonConst([j]); prev = X.CloneView(list(range(j)))
        old = Xj.CloneCopy()
        pr = Xj.MvTransMv(1.0, prev)
        Xj.MvTimesMatAddMv(-1.0, prev, pr, 1.0)
        d = Xj.MvDot(Xj); Xj.MvScale(1.0 / np.sqrt(d[0]))
    print("  findBasis steps ok", ctx.findbasis_stats())
ng, rp, ci, va = tb.galeri("Laplace3D", 20, 20, 20)
op = tb.CrsOperator(ctx, ng, 0, ng, rp, ci, va)
x = np.asfortranarray(rng.uniform(-1, 1, (ng, 4)))
Xs = tb.MultiVec.from_numpy(ctx, x); Ys = tb.MultiVec(ctx, ng, 4)
op.Apply(Xs, Ys)
print("spmm", rel(Ys.to_numpy(), oracle.spmm(rp, ci.astype(np.int32), va, x)))
ctx.sync()
print("done")
