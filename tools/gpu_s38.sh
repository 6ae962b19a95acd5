#!/bin/bash
# session 38 (one GPU): the per-launch prefetch distance (auto) against both forced values; whole GPU suite; bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/s38_spmm_pf_auto.txt; : > $out
for v in auto 1 2; do
  echo "## B200_WIN_PF=$v" >> $out
  if [ $v = auto ]; then unset B200_WIN_PF; else export B200_WIN_PF=$v; fi
  python tools/spmm_bench.py Laplace3D 100 4 >> $out 2>&1
  python tools/spmm_bench.py Laplace3D 200 4 cold >> $out 2>&1
  python tools/spmm_bench.py Recirc2D 2000 16 cold >> $out 2>&1
done
unset B200_WIN_PF
cat $out
python -m pytest tests -m gpu -x -q > gpurun_out/s38_pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -n 2 gpurun_out/s38_pytest_gpu.txt
python bench.py > gpurun_out/s38_bench_n1.json 2> gpurun_out/s38_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/s38_bench_n1.json') if l.startswith('{')][-1])
print('config2', d['value'], 'e2e', d['e2e']['value'], 'spmm', d['roofline']['kernels']['spmm'])
n=d['north_star_8m']; print('8m', n['value'], n['roofline']['kernels']['spmm'])
c=d['c5_cycle']; print('c5', c['value'], c['roofline']['kernels']['spmm'])
PY
