#!/bin/bash
# session 41: final state of the round on one GPU -- whole GPU suite, smoke, default bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/s41_pytest_gpu.txt 2>&1; echo "suite rc=$?"; tail -n 1 gpurun_out/s41_pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s41_smoke.txt 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/s41_bench_n1.json 2> gpurun_out/s41_bench_n1.err; echo "bench rc=$?"
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/s41_bench_n1.json') if l.startswith('{')][-1])
print('config2', d['value'], 'e2e', d['e2e']['value'], 'roofline', d['roofline']['frac'], 'solve', d['roofline']['solve'])
print('8m', d['north_star_8m']['value'], 'c5', d['c5_cycle']['value'], 'cpu', d['cpu_baseline']['value'])
PY
