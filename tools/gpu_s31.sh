#!/bin/bash
# final kernel table + ncu of the Gram kernel after the equal-width-block change
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 300 python tools/kbench.py 1000000 4 > gpurun_out/r2_kbench_s31.txt 2>&1
grep -v "^{" gpurun_out/r2_kbench_s31.txt | tail -30
python tools/ncu_kernels.py > gpurun_out/ncu_k_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:gram_tma -s 1 -c 1 -o gpurun_out/r2_gram_s31 -f python tools/ncu_kernels.py > gpurun_out/ncu_k.log 2>&1
tail -2 gpurun_out/ncu_k.log
