"""BASELINE config 5 (Recirc2D 2000^2, block size 16, restart 100) beyond its first restart cycle: the native residual
norms at the end of every cycle, to document how restarted Block GMRES progresses on this convection-dominated operator
(DESIGN.md 6b).  usage: python tools/c5_progress.py [cycles]"""
import sys, json
import numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb
from bench import WORKLOADS

cycles = int(sys.argv[1]) if len(sys.argv) > 1 else 15
name, dims, b, m, ortho, tol, _ = WORKLOADS["recirc2d_2000_b16"]
ctx = tb.Context(0)
op, n, lo, hi = tb.dist.local_operator(ctx, name, dims)
ctx.set_seed(1)
Xe = tb.dist.local_multivec(ctx, n, lo, hi, b, zero=False); Xe.MvRandom()
B = tb.dist.local_multivec(ctx, n, lo, hi, b); op.Apply(Xe, B)
X = tb.dist.local_multivec(ctx, n, lo, hi, b)
mgr = tb.BlockGmresSolMgr(op, block_size=b, num_blocks=m, max_iters=m * cycles, max_restarts=100000, tol=tol, ortho=ortho)
r = mgr.solve(X, B, history=m * cycles + 8)
h = np.array(r["native_resid_history"])
h0 = h[0]
print(f"Recirc2D 2000^2 b={b} m={m}: {r['num_iters']} block iterations in {r['solve_seconds']:.1f} s, converged={r['converged']}")
for c in range(0, len(h), m):
    rel = h[c] / h0
    print(f"  after {c:5d} block iterations: native residual / initial  max {rel.max():.3e}  min {rel.min():.3e}")
rel = h[-1] / h0
print(f"  after {len(h) - 1:5d} block iterations: native residual / initial  max {rel.max():.3e}  min {rel.min():.3e}")
print("true relative residuals:", " ".join(f"{v:.2e}" for v in r["true_rel_resid"]))
