#!/bin/bash
# GPU session 8: why is the fused update + Gram kernel slow?  knob sweep + one ncu capture
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/r2_ug_sweep_s8.txt; : > $out
run() { echo "## $*" >> $out; env "$@" 2>&1 | tail -1 >> $out; }
export B200_SPEC_GRAM=2
run python tools/ncu_ug.py 1000000 200 4
run B200_UG_DBG=1 python tools/ncu_ug.py 1000000 200 4
run B200_UG_DBG=2 python tools/ncu_ug.py 1000000 200 4
run B200_UG_DBG=3 python tools/ncu_ug.py 1000000 200 4
run B200_UG_ROWS=28 python tools/ncu_ug.py 1000000 200 4
run B200_UG_ROWS=28 B200_UG_DBG=3 python tools/ncu_ug.py 1000000 200 4
run python tools/ncu_ug.py 1000000 16 4
run B200_UG_DBG=3 python tools/ncu_ug.py 1000000 16 4
run B200_UG_ROWS=60 python tools/ncu_ug.py 1000000 16 4
run B200_UG_ROWS=60 B200_UG_DBG=3 python tools/ncu_ug.py 1000000 16 4
cat $out
python tools/ncu_ug.py 1000000 200 4 2 > gpurun_out/ncu_ug_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:update_gram -s 3 -c 1 -o gpurun_out/r2_ug_s8 -f python tools/ncu_ug.py 1000000 200 4 2 > gpurun_out/ncu_ug.log 2>&1
tail -3 gpurun_out/ncu_ug.log
