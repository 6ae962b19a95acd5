"""Time to solution of BASELINE config 2 (Laplace3D 100^3, b = 4, 50 blocks, DGKS, 1e-8) without and with the Chebyshev right
preconditioner.  usage: python tools/prec_demo.py [degrees...]"""
import sys, json
sys.path.insert(0, ".")
import trilinos_b200 as tb

degrees = [int(a) for a in sys.argv[1:]] or [0, 2, 4, 8, 16]
ctx = tb.Context(0)
n, rp, ci, va = tb.galeri("Laplace3D", 100, 100, 100)
op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
ctx.set_seed(1)
Xe = tb.MultiVec(ctx, n, 4, zero=False); Xe.MvRandom()
B = tb.MultiVec(ctx, n, 4); op.Apply(Xe, B)
for d in degrees:
    P = tb.Chebyshev(op, degree=d) if d > 0 else None
    mgr = tb.BlockGmresSolMgr(op, block_size=4, num_blocks=50, max_iters=-1, max_restarts=1000, tol=1e-8, ortho="DGKS", right_prec=P)
    best = None
    for rep in range(3):
        X = tb.MultiVec(ctx, n, 4)
        r = mgr.solve(X, B)
        best = r if best is None or r["solve_seconds"] < best["solve_seconds"] else best
    print(json.dumps(dict(cheby_degree=d, lambda_max=P.lambdas()[0] if P else None, iters=best["num_iters"], converged=best["converged"],
                          solve_ms=round(best["solve_seconds"] * 1e3, 2), true_rel_resid=max(best["true_rel_resid"]))), flush=True)
