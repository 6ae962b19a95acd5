#!/bin/bash
# session 37 (one GPU): descriptor prefetch distance of the window SpMM's producer warp, 2 tiles ahead (default build)
# vs 1 (build/v_pf1, -DB200_WIN_PF=1); parity tests of the apply with the default build
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/s37_spmm_pf.txt; : > $out
python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "apply or spmm" > gpurun_out/s37_pytest_apply.txt 2>&1; echo "apply tests rc=$?" | tee -a $out
tail -n 1 gpurun_out/s37_pytest_apply.txt | tee -a $out
cp trilinos_b200/lib/libb200belos.so /tmp/lib_pf2.so
for v in pf2 pf1 pf2 pf1; do
  if [ $v = pf2 ]; then cp /tmp/lib_pf2.so trilinos_b200/lib/libb200belos.so; else cp build/v_pf1/lib/libb200belos.so trilinos_b200/lib/libb200belos.so; fi
  echo "## $v" >> $out
  python tools/spmm_bench.py Laplace3D 100 4 >> $out 2>&1
  python tools/spmm_bench.py Laplace3D 200 4 cold >> $out 2>&1
  python tools/spmm_bench.py Brick3D 150 8 cold >> $out 2>&1
  python tools/spmm_bench.py Recirc2D 2000 16 cold >> $out 2>&1
done
cp /tmp/lib_pf2.so trilinos_b200/lib/libb200belos.so
cat $out
