#!/bin/bash
# GPU session 2: window-SpMM geometry sweep, Gram narrow-block mapping (kbench), quick regression tests
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/r2_sweep_s2.txt; : > $out
run() { echo "## $*" >> $out; env "$@" 2>&1 | grep -E "spmm|Error|error" >> $out; }
for m in "Laplace3D 100 4" "Laplace3D 200 4"; do
  run B200_SPMM_WIN=1 python tools/spmm_bench.py $m cold
  run B200_SPMM_WIN=1 B200_WIN_CTAS=1 python tools/spmm_bench.py $m cold
  run B200_SPMM_WIN=1 B200_WIN_R=128 python tools/spmm_bench.py $m cold
  run B200_SPMM_WIN=1 B200_WIN_R=128 B200_WIN_CTAS=2 python tools/spmm_bench.py $m cold
  run B200_SPMM_WIN=1 B200_WIN_R=128 B200_WIN_STAGES=2 python tools/spmm_bench.py $m cold
done
run B200_SPMM_WIN=1 python tools/spmm_bench.py Laplace3D 100 1 cold
run B200_SPMM_WIN=1 python tools/spmm_bench.py Laplace2D 1000 4 cold
run B200_SPMM_WIN=1 python tools/spmm_bench.py Recirc2D 2000 8 cold
run B200_SPMM_WIN=0 python tools/spmm_bench.py Recirc2D 2000 8 cold
cat $out
timeout 600 python tools/kbench.py 1000000 4 > gpurun_out/r2_kbench_s2.txt 2>&1; grep -E "trans_mv|round trip" gpurun_out/r2_kbench_s2.txt
timeout 900 python -m pytest tests/test_gpu_kernels.py tests/test_gpu_solver.py -m gpu -x -q 2>&1 | tail -5
