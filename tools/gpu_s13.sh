#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
export B200_SPEC_GRAM=2
python tools/ncu_ug.py 1000000 200 4 2 > gpurun_out/ncu_ug_plain.log 2>&1 && \
ncu --set full --clock-control none --import-source on -k regex:update_gram -s 3 -c 1 -o gpurun_out/r2_ug200_s18 -f python tools/ncu_ug.py 1000000 200 4 2 > gpurun_out/ncu_ug200.log 2>&1
tail -2 gpurun_out/ncu_ug200.log
