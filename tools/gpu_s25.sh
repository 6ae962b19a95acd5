#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
cp trilinos_b200/lib/libb200belos.so /tmp/lib_full.so
cp build/v_trace/lib/libb200belos.so trilinos_b200/lib/libb200belos.so
export B200_SPEC_GRAM=2
for p in 200 100 32; do
  B200_UG_TRACE_FILE=gpurun_out/ug_trace_p$p.txt python tools/ncu_ug.py 1000000 $p 4 2 | tail -1
  python tools/ug_trace.py gpurun_out/ug_trace_p$p.txt 20 10
done
cp /tmp/lib_full.so trilinos_b200/lib/libb200belos.so
