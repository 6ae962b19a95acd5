#!/bin/bash
# GPU session 1 (round 2): tests, SpMM A/B (window vs gather kernel), bench, ncu of the window kernel
set -x
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_s1.txt 2>&1
tail -15 gpurun_out/r2_pytest_s1.txt
: > gpurun_out/r2_spmm_ab_s1.txt
for m in "Laplace3D 100 4" "Laplace3D 200 4" "Laplace3D 100 1" "Brick3D 150 8" "Recirc2D 2000 16" "Laplace2D 1000 4"; do
  for w in 0 1; do
    B200_SPMM_WIN=$w timeout 300 python tools/spmm_bench.py $m >> gpurun_out/r2_spmm_ab_s1.txt 2>&1
  done
done
cat gpurun_out/r2_spmm_ab_s1.txt
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_s1.json 2> gpurun_out/r2_bench_s1.err
tail -c 3000 gpurun_out/r2_bench_s1.err
cat gpurun_out/r2_bench_s1.json
timeout 600 ncu --set full --clock-control none --import-source on -k regex:spmm_win -c 3 -o gpurun_out/r2_spmm_win_s1 \
  python tools/spmm_bench.py Laplace3D 100 4 cold > gpurun_out/r2_ncu_s1.log 2>&1
tail -3 gpurun_out/r2_ncu_s1.log
