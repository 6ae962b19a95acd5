#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused_update_gram or times_mat_add_mv or fused_update_dots" 2>&1 | tail -30
out=gpurun_out/r2_ug_sweep_s12.txt; : > $out
run() { echo "## $*" >> $out; env "$@" 2>&1 | tail -1 >> $out; }
export B200_SPEC_GRAM=2
for p in 200 152 100 64 32 16; do
run python tools/ncu_ug.py 1000000 $p 4
done
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 100 4
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 32 4
run B200_UG_ROWS=124 python tools/ncu_ug.py 1000000 32 4
run python tools/ncu_ug.py 1000000 200 8
run python tools/ncu_ug.py 125000 200 4
run python tools/ncu_ug.py 125000 100 4
cat $out
out2=gpurun_out/r2_ug_sweep_s12_dmma.txt; : > $out2
for p in 200 100 32; do echo "## DMMA update p=$p" >> $out2; B200_UG_DMMA_UPDATE=1 python tools/ncu_ug.py 1000000 $p 4 2>&1 | tail -1 >> $out2; done
cat $out2
