"""Launches each hot kernel once at BASELINE config-2 shapes (n = 1M, b = 4, basis width p) — the command
profiled with `ncu --set full`.  usage: python tools/ncu_kernels.py [n] [p] [q]"""
import sys
import numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
p = int(sys.argv[2]) if len(sys.argv) > 2 else 100
q = int(sys.argv[3]) if len(sys.argv) > 3 else 4
ctx = tb.Context(0)
ctx.set_seed(1)
Q = tb.MultiVec(ctx, n, p, zero=False); Q.MvRandom()
X = tb.MultiVec(ctx, n, q, zero=False); X.MvRandom()
nx = round(n ** (1 / 3))
ng, rp, ci, va = tb.galeri("Laplace3D", nx, nx, nx)
op = tb.CrsOperator(ctx, ng, 0, ng, rp, ci, va)
Y = X.Clone(q)
Bm = np.random.default_rng(0).uniform(-1, 1, (p, q)) * 1e-3
x1 = X.CloneViewNonConst([0]); Q3 = Q.CloneView([0, 1, 2])
for rep in range(2):       # second round = warm instruction caches / pool; ncu -s skips the first
    C = X.MvTransMv(1.0, Q)
    X.MvTimesMatAddMv(-1.0, Q, Bm, 1.0)
    op.Apply(X, Y)
    c3 = x1.MvTransMv(1.0, Q3)
    x1.MvTimesMatAddMv(-1.0, Q3, Bm[:3, :1], 1.0)
    ctx.set_fusion(False); d = x1.MvDot(x1); ctx.set_fusion(True)
    Z = X.CloneCopy()
ctx.sync()
print("ok", C.shape)
