"""Solve one Galeri config on the GPU: python tools/solve_demo.py Laplace3D 100 4 50 DGKS 1e-8 [timers|-] [reps]"""
import sys, time, json
import numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb
import oracle

name = sys.argv[1] if len(sys.argv) > 1 else "Laplace3D"
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 100
b = int(sys.argv[3]) if len(sys.argv) > 3 else 4
m = int(sys.argv[4]) if len(sys.argv) > 4 else 50
ortho = sys.argv[5] if len(sys.argv) > 5 else "DGKS"
tol = float(sys.argv[6]) if len(sys.argv) > 6 else 1e-8
timers = len(sys.argv) > 7 and sys.argv[7] == "timers"
reps = int(sys.argv[8]) if len(sys.argv) > 8 else 2
dims = (nx, nx, nx) if name in ("Laplace3D", "Brick3D") else (nx, nx, 1)
ctx = tb.Context(0)
n, rp, ci, va = tb.galeri(name, *dims)
op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
xe, _ = oracle.random(1, n, b)
B = tb.MultiVec(ctx, n, b)
Xe = tb.MultiVec.from_numpy(ctx, xe)
op.Apply(Xe, B)
for rep in range(reps):
    X = tb.MultiVec(ctx, n, b)
    mgr = tb.BlockGmresSolMgr(op, block_size=b, num_blocks=m, max_iters=-1, max_restarts=1000, tol=tol, ortho=ortho, show_timers=timers)
    l0, s0 = ctx.counters()
    r = mgr.solve(X, B)
    l1, s1 = ctx.counters()
    r["launches"] = l1 - l0; r["syncs"] = s1 - s0
    r["iters_per_s"] = r["num_iters"] / r["solve_seconds"]
    print(json.dumps(r))
err = X.to_numpy() - xe
print("max |x - x_exact|:", np.abs(err).max())
