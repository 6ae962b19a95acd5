"""Operator-apply micro-benchmark (CUDA events on the compute stream, L2 flushed between launches).
usage: python tools/spmm_bench.py [matrix] [nx] [k]   e.g.  Laplace3D 100 4 | Brick3D 150 8 | Recirc2D 2000 16"""
import json
import os
import sys
sys.path.insert(0, ".")
import trilinos_b200 as tb

PEAK = float(json.load(open("MEASURED_PEAKS.json"))["hbm_gbs"]) if os.path.exists("MEASURED_PEAKS.json") else 6650.0
KERNEL = "gather" if os.environ.get("B200_SPMM_WIN") == "0" else "window"

name = sys.argv[1] if len(sys.argv) > 1 else "Laplace3D"
nx = int(sys.argv[2]) if len(sys.argv) > 2 else 100
k = int(sys.argv[3]) if len(sys.argv) > 3 else 4
dims = (nx, nx) if name.endswith("2D") else (nx, nx, nx)
ctx = tb.Context(0)
ctx.set_seed(1)
n, rp, ci, va = tb.galeri(name, *dims)
op = tb.CrsOperator(ctx, n, 0, n, rp, ci, va)
X = tb.MultiVec(ctx, n, k, zero=False); X.MvRandom()
Y = X.Clone(k)
modes = (True,) if (len(sys.argv) > 4 and sys.argv[4] == "cold") else (True, False)
for flush in modes:
    for _ in range(3): op.Apply(X, Y)
    ctx.sync(); ctx.timing_read(reset=True); ctx.timing_enable(True)
    for _ in range(20):
        if flush: ctx.flush_l2()
        op.Apply(X, Y)
    ctx.timing_enable(False)
    t = ctx.timing_read(reset=True)["spmm"]
    us = t["ms"] / t["calls"] * 1e3
    gbs = t["bytes"] / (t["ms"] * 1e-3) / 1e9
    print(f"spmm[{KERNEL}] {name} {nx} k={k} nnz={len(va)} {'cold' if flush else 'L2-warm'}: {us:8.1f} us  {gbs:8.1f} GB/s  frac={gbs/PEAK:.3f}", flush=True)
