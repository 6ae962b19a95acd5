"""Driver for the ncu capture of the 16-column Gram / update kernels (n = 1M, p = 200, q = 16).
usage: [B200_GRAM_DMMA=0 B200_UPDATE_DMMA=0] python tools/ncu_b16.py      (SIMT or DMMA kernels)"""
import sys
import numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb

n, p, q = 1_000_000, 200, 16
ctx = tb.Context(0)
ctx.set_seed(1)
Q = tb.MultiVec(ctx, n, p, zero=False); Q.MvRandom()
X = tb.MultiVec(ctx, n, q, zero=False); X.MvRandom()
Bm = np.random.default_rng(0).uniform(-1, 1, (p, q)) * 1e-3
for _ in range(2):
    X.MvTransMv(1.0, Q)
    X.MvTimesMatAddMv(-1.0, Q, Bm, 1.0)
ctx.sync()
