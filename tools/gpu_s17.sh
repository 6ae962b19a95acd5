#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/r2_ug_sweep_s17.txt; : > $out
run() { echo "## $*" >> $out; env "$@" 2>&1 | tail -1 >> $out; }
export B200_SPEC_GRAM=2
run B200_UG_ROWS=28 python tools/ncu_ug.py 1000000 200 4
run B200_UG_ROWS=28 B200_UG_BUFS=3 python tools/ncu_ug.py 1000000 200 4
run B200_UG_ROWS=28 python tools/ncu_ug.py 1000000 100 4
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 100 4
run B200_UG_ROWS=60 B200_UG_BUFS=2 python tools/ncu_ug.py 1000000 100 4
run B200_UG_ROWS=28 python tools/ncu_ug.py 1000000 64 4
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 64 4
run B200_UG_ROWS=124 python tools/ncu_ug.py 1000000 64 4
cat $out
