#!/bin/bash
# session 40, N = $1 ranks: bench.py of the final state (config 2, 8M rows, config-5 cycle)
N=${1:-4}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout ${B200_RUN_LIMIT:-280} python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29640 \
  bench.py --gpus $N --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/s40_bench_n$N.json 2> gpurun_out/s40_bench_n$N.err; echo "bench rc=$?"
python - $N <<'PY'
import json, sys
d=json.loads([l for l in open('gpurun_out/s40_bench_n%s.json'%sys.argv[1]) if l.startswith('{')][-1])
print('config2', round(d['value'],1), 'parity', d.get('parity_check',{}).get('ok'))
for key in ('north_star_8m','c5_cycle'):
    n=d.get(key)
    if n: print(key, round(n['value'],1), 'spmm us/call', round(1000*n['roofline']['kernels']['spmm']['ms']/n['roofline']['kernels']['spmm']['calls'],1))
PY
