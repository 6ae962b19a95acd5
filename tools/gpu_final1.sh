#!/bin/bash
# final single-GPU session of the round: smoke, full GPU test-suite, bench (with the CPU baseline), reference arm
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/r2_smoke_final.txt 2>&1; tail -2 gpurun_out/r2_smoke_final.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_final.txt 2>&1
tail -4 gpurun_out/r2_pytest_final.txt
timeout 900 python bench.py --impl reference > gpurun_out/r2_bench_reference_final.json 2> gpurun_out/r2_bench_reference_final.err
tail -c 400 gpurun_out/r2_bench_reference_final.json
timeout 1200 python bench.py > gpurun_out/r2_bench_final.json 2> gpurun_out/r2_bench_final.err
tail -c 600 gpurun_out/r2_bench_final.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_final.json') if l.startswith('{')][-1])
print('value',d['value'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'],'syncs',d['host_syncs'],'cpu',d['cpu_baseline']['value'])
print({k:(v['calls'],round(v['ms'],1),v['frac']) for k,v in d['roofline']['kernels'].items()})
print(d['roofline']['kernel'], d['roofline']['frac'], d['roofline']['solve'])
for k in ('north_star_8m','c5_cycle'):
    v=d[k]; print(k, v['value'], {kk:(vv['calls'],round(vv['ms'],1),vv['frac']) for kk,vv in v['roofline']['kernels'].items()})
PY
