#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_s32.txt 2>&1
tail -4 gpurun_out/r2_pytest_s32.txt
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_s32.json 2> gpurun_out/r2_bench_s32.err
tail -c 300 gpurun_out/r2_bench_s32.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_s32.json') if l.startswith('{')][-1])
print('value',d['value'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'],'syncs',d['host_syncs'])
print({k:(v['calls'],round(v['ms'],1),v['frac']) for k,v in d['roofline']['kernels'].items()})
for k in ('north_star_8m','c5_cycle'):
    v=d[k]; print(k, v['value'])
PY
