"""One restart cycle (Num Blocks block iterations) of a bench workload — the command profiled by ncu.
usage: python tools/profile_cycle.py [workload] [cycles]"""
import sys, json
sys.path.insert(0, ".")
import trilinos_b200 as tb
from bench import WORKLOADS

wl = sys.argv[1] if len(sys.argv) > 1 else "laplace3d_100_b4"
cycles = int(sys.argv[2]) if len(sys.argv) > 2 else 1
name, dims, b, m, ortho, tol, _ = WORKLOADS[wl]
ctx = tb.Context(0)
op, n, lo, hi = tb.dist.local_operator(ctx, name, dims)
ctx.set_seed(1)
Xe = tb.dist.local_multivec(ctx, n, lo, hi, b, zero=False); Xe.MvRandom()
B = tb.dist.local_multivec(ctx, n, lo, hi, b); op.Apply(Xe, B)
X = tb.dist.local_multivec(ctx, n, lo, hi, b)
mgr = tb.BlockGmresSolMgr(op, block_size=b, num_blocks=m, max_iters=m * cycles, max_restarts=100000, tol=tol, ortho=ortho)
import os
timing = os.environ.get("B200_CYCLE_TIMING") == "1"   # per-kernel-class CUDA-event times (adds event overhead)
if timing:
    ctx.timing_read(reset=True); ctx.timing_enable(True)
l0, s0 = ctx.counters()
r = mgr.solve(X, B)
l1, s1 = ctx.counters()
out = dict(workload=wl, iters=r["num_iters"], seconds=r["solve_seconds"], launches=l1 - l0, syncs=s1 - s0)
if timing:
    ctx.timing_enable(False)
    out["classes"] = {k: dict(calls=v["calls"], ms=round(v["ms"], 2), GBps=round(v["bytes"] / max(v["ms"], 1e-9) / 1e6, 1))
                      for k, v in ctx.timing_read(reset=True).items() if v["calls"]}
print(json.dumps(out))
