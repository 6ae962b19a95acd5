#!/bin/bash
# GPU session 10: full GPU test-suite + bench with the fused update + Gram kernel
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_s10.txt 2>&1
tail -8 gpurun_out/r2_pytest_s10.txt
for mode in 1 0; do
B200_SPEC_GRAM=$mode timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_s10_spec$mode.json 2> gpurun_out/r2_bench_s10_spec$mode.err
tail -c 600 gpurun_out/r2_bench_s10_spec$mode.err
python - $mode <<'PY'
import json,sys
d=json.loads([l for l in open('gpurun_out/r2_bench_s10_spec%s.json'%sys.argv[1]) if l.startswith('{')][-1])
print('SPEC',sys.argv[1],'value',d['value'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'],'syncs',d['host_syncs'])
print({k:(v['calls'],round(v['ms'],1),v['frac']) for k,v in d['roofline']['kernels'].items()})
for k in ('north_star_8m','c5_cycle'):
    v=d[k]; print(k, v['value'], {kk:(vv['calls'],round(vv['ms'],1),vv['frac']) for kk,vv in v['roofline']['kernels'].items()})
PY
done
