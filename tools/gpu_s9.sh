#!/bin/bash
# GPU session 9: fused update + Gram kernel after the predication fix -- tests, knob sweep
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused_update_gram or times_mat_add_mv or fused_update_dots" > gpurun_out/r2_pytest_ug_s9.txt 2>&1
tail -5 gpurun_out/r2_pytest_ug_s9.txt
out=gpurun_out/r2_ug_sweep_s9.txt; : > $out
run() { echo "## $*" >> $out; env "$@" 2>&1 | tail -1 >> $out; }
export B200_SPEC_GRAM=2
for p in 200 100 32 16; do
run python tools/ncu_ug.py 1000000 $p 4
run B200_UG_DBG=1 python tools/ncu_ug.py 1000000 $p 4
run B200_UG_DBG=2 python tools/ncu_ug.py 1000000 $p 4
run B200_UG_DBG=3 python tools/ncu_ug.py 1000000 $p 4
done
run B200_UG_ROWS=28 python tools/ncu_ug.py 1000000 200 4
run B200_UG_ROWS=28 B200_UG_DBG=3 python tools/ncu_ug.py 1000000 200 4
run B200_UG_ROWS=28 B200_UG_BUFS=2 python tools/ncu_ug.py 1000000 200 4
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 100 4
run B200_UG_ROWS=28 python tools/ncu_ug.py 1000000 100 4
run B200_UG_ROWS=124 python tools/ncu_ug.py 1000000 32 4
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 32 4
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 16 4
run python tools/ncu_ug.py 1000000 200 8
run python tools/ncu_ug.py 1000000 100 8
cat $out
