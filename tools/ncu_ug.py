"""Fused update + Gram kernel at one shape: timing by CUDA events (default) or a few launches for ncu.
usage: python tools/ncu_ug.py [n] [p] [q] [reps]"""
import sys
import numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
p = int(sys.argv[2]) if len(sys.argv) > 2 else 200
q = int(sys.argv[3]) if len(sys.argv) > 3 else 4
reps = int(sys.argv[4]) if len(sys.argv) > 4 else 10
ctx = tb.Context(0)
ctx.set_seed(1)
Q = tb.MultiVec(ctx, n, p, zero=False); Q.MvRandom()
X = tb.MultiVec(ctx, n, q, zero=False); X.MvRandom()
Bm = np.random.default_rng(0).uniform(-1, 1, (p, q)) * 1e-3
for _ in range(3):
    X.MvTimesMatAddMv(-1.0, Q, Bm, 1.0)
ctx.sync(); ctx.timing_read(reset=True); ctx.timing_enable(True)
for _ in range(reps):
    ctx.flush_l2()
    X.MvTimesMatAddMv(-1.0, Q, Bm, 1.0)
ctx.timing_enable(False)
tr = ctx.timing_read(reset=True)
t = tr["times_mat_add_mv_with_next_trans_mv"] if tr["times_mat_add_mv_with_next_trans_mv"]["calls"] else tr["times_mat_add_mv"]
print(f"n={n} p={p} q={q}: {t['ms'] / t['calls'] * 1e3:.1f} us per update, spec launches/hits {ctx.spec_gram_stats()}")
