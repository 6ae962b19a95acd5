#!/bin/bash
# role isolation of the fused update + Gram kernel: variants built with -DB200_UG_NO_GRAM / -DB200_UG_NO_UPDATE
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/r2_ug_roles_s19.txt; : > $out
export B200_SPEC_GRAM=2
cp trilinos_b200/lib/libb200belos.so /tmp/lib_full.so
for v in full v_nogram v_noupd v_skel; do
  if [ $v = full ]; then cp /tmp/lib_full.so trilinos_b200/lib/libb200belos.so; else cp build/$v/lib/libb200belos.so trilinos_b200/lib/libb200belos.so; fi
  for p in 200 100 32; do
    echo "## $v p=$p simt-update" >> $out; python tools/ncu_ug.py 1000000 $p 4 2>&1 | tail -1 >> $out
    echo "## $v p=$p dmma-update" >> $out; B200_UG_DMMA_UPDATE=1 python tools/ncu_ug.py 1000000 $p 4 2>&1 | tail -1 >> $out
  done
done
cp /tmp/lib_full.so trilinos_b200/lib/libb200belos.so
cat $out
