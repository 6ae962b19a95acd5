#!/bin/bash
# GPU session 4: full tests, SpMM benches incl. K-split (Recirc2D k=16), bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_s4.txt 2>&1
tail -15 gpurun_out/r2_pytest_s4.txt
out=gpurun_out/r2_sweep_s4.txt; : > $out
run() { echo "## $*" >> $out; env "$@" 2>&1 | grep -E "spmm|Error|error" >> $out; }
for m in "Laplace3D 100 4" "Laplace3D 200 4" "Brick3D 150 8" "Recirc2D 2000 16" "Laplace3D 100 16" "Laplace2D 1000 8"; do
  run B200_SPMM_WIN=1 python tools/spmm_bench.py $m cold
done
cat $out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_s4.json 2> gpurun_out/r2_bench_s4.err
tail -c 2000 gpurun_out/r2_bench_s4.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_s4.json') if l.startswith('{')][-1])
print('value',d['value'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'],'syncs',d['host_syncs'])
print({k:(v['calls'],v['frac']) for k,v in d['roofline']['kernels'].items()})
for k in ('north_star_8m','c5_cycle'):
    v=d[k]; print(k, v['value'], {kk:(vv['calls'],vv['frac']) for kk,vv in v['roofline']['kernels'].items()})
PY
