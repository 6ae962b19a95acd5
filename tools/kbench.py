"""Kernel micro-benchmark at BASELINE config sizes (CUDA events on the compute stream).
usage: python tools/kbench.py [n] [b]"""
import sys, json, os
import numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb
import oracle

n = int(sys.argv[1]) if len(sys.argv) > 1 else 1_000_000
b = int(sys.argv[2]) if len(sys.argv) > 2 else 4
PEAK = float(json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]) if os.path.exists("MEASURED_PEAKS.json") else 6650.0
ctx = tb.Context(0)
ctx.set_seed(1)
PMAX = 200
Q = tb.MultiVec(ctx, n, PMAX, zero=False); Q.MvRandom()
X = tb.MultiVec(ctx, n, b, zero=False); X.MvRandom()
nx = round(n ** (1 / 3))
ng, rp, ci, va = oracle.galeri_csr("Laplace3D", nx, nx, nx)
assert ng == n
op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
Y = X.Clone(b)
res = {}
def timed(name, fn, reps=20, flush=True):
    for _ in range(3): fn()
    ctx.sync(); ctx.timing_read(reset=True); ctx.timing_enable(True)
    for _ in range(reps):
        if flush: ctx.flush_l2()
        fn()
    ctx.timing_enable(False)
    t = ctx.timing_read(reset=True)
    for k, v in t.items():
        if v["calls"]:
            gbs = v["bytes"] / (v["ms"] * 1e-3) / 1e9
            res[name] = dict(cls=k, ms=v["ms"] / v["calls"], GBs=gbs)
            print(f"{name:34s} {k:18s} {v['ms']/v['calls']*1e3:9.1f} us  {gbs:8.1f} GB/s  frac_of_peak={gbs/PEAK:.3f}", flush=True)
for p in (4, 16, 32, 64, 100, 200):
    Qp = Q.CloneView(list(range(p)))
    timed(f"trans_mv p={p} q={b}", lambda: X.MvTransMv(1.0, Qp))
    Bm = np.random.default_rng(0).uniform(-1, 1, (p, b)) * 1e-3
    timed(f"times_mat_add_mv p={p} q={b}", lambda: X.MvTimesMatAddMv(-1.0, Qp, Bm, 1.0))
x1 = X.CloneView([0])
for p in (1, 3):
    Qp = Q.CloneView(list(range(p)))
    timed(f"trans_mv p={p} q=1 (L2)", lambda: x1.MvTransMv(1.0, Qp), flush=False)
    Bs = np.random.default_rng(0).uniform(-1, 1, (p, 1)) * 1e-3
    timed(f"times_mat_add_mv p={p} q=1 (L2)", lambda: x1.MvTimesMatAddMv(-1.0, Qp, Bs, 1.0), flush=False)
timed(f"dot k={b} (L2)", lambda: X.MvDot(X), flush=False)
timed(f"dot k=1 (L2)", lambda: x1.MvDot(x1), flush=False)
timed(f"norm k={b} (L2)", lambda: X.MvNorm(), flush=False)
timed(f"spmm k={b}", lambda: op.Apply(X, Y))
timed(f"spmm k={b} (L2 warm)", lambda: op.Apply(X, Y), flush=False)
timed(f"clonecopy k={b}", lambda: X.CloneCopy(), flush=False)
# host-visible latency of a reduction (launch + kernel + mailbox)
import time
ctx.sync()
t0 = time.perf_counter()
for _ in range(200): x1.MvDot(x1)
t1 = time.perf_counter()
print(f"dot k=1 round trip: {(t1-t0)/200*1e6:.1f} us")
for p in (4, 32):
    Qp = Q.CloneView(list(range(p)))
    X.MvTransMv(1.0, Qp); ctx.sync()
    t0 = time.perf_counter()
    for _ in range(200): X.MvTransMv(1.0, Qp)
    t1 = time.perf_counter()
    print(f"trans_mv p={p} q={b} host round trip (L2 warm): {(t1-t0)/200*1e6:.1f} us")
Q3 = Q.CloneView([0, 1, 2])
t0 = time.perf_counter()
for _ in range(200): x1.MvTransMv(1.0, Q3)
t1 = time.perf_counter()
print(f"trans_mv p=3 q=1 host round trip: {(t1-t0)/200*1e6:.1f} us")
t0 = time.perf_counter()
for _ in range(200): X.MvScale(1.0000001)
ctx.sync()
t1 = time.perf_counter()
print(f"async scale launch rate: {(t1-t0)/200*1e6:.1f} us per op")
print(json.dumps(res))
