#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "trans_mv" 2>&1 | tail -3
B200_SPEC_GRAM=0 timeout 300 python tools/kbench.py 1000000 4 2>&1 | grep -E "trans_mv p=" > gpurun_out/r2_kbench_gram_s22.txt
cat gpurun_out/r2_kbench_gram_s22.txt
python - <<'PY'
import sys, numpy as np
sys.path.insert(0, ".")
import trilinos_b200 as tb
ctx = tb.Context(0); ctx.set_seed(1)
n = 1000000
Q = tb.MultiVec(ctx, n, 200, zero=False); Q.MvRandom()
X = tb.MultiVec(ctx, n, 4, zero=False); X.MvRandom()
for p in (36, 68, 100, 132, 164, 196):
    Qp = Q.CloneView(list(range(p)))
    for _ in range(3): X.MvTransMv(1.0, Qp)
    ctx.sync(); ctx.timing_read(reset=True); ctx.timing_enable(True)
    for _ in range(20):
        ctx.flush_l2(); X.MvTransMv(1.0, Qp)
    ctx.timing_enable(False)
    t = ctx.timing_read(reset=True)["trans_mv"]
    us = t["ms"] / t["calls"] * 1e3
    print(f"trans_mv p={p} q=4: {us:.1f} us  frac {8.0*n*(p+4)/us/1e3/6547.8:.3f}")
PY
