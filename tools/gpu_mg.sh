#!/bin/bash
# multi-GPU session: N = $1 ranks.  Parity worker (tests/_nccl_worker.py through pytest), then bench.py under torchrun.
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
nvidia-smi -L | head -8
timeout 1200 python -m pytest tests/test_multigpu_nccl.py -m gpu -x -q > gpurun_out/r2_pytest_mg_n$N.txt 2>&1
tail -25 gpurun_out/r2_pytest_mg_n$N.txt
timeout 1200 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 \
  bench.py --gpus $N --steps 5 --warmup 3 > gpurun_out/r2_bench_n$N.json 2> gpurun_out/r2_bench_n$N.err
tail -c 1500 gpurun_out/r2_bench_n$N.err
python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_n$N.json') if l.startswith('{')][-1])
print('value',d['value'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'],'syncs',d['host_syncs'], d['parity_check']['max_rel_err'])
print({k:(v['calls'],v['frac']) for k,v in d['roofline']['kernels'].items()})
for k in ('north_star_8m','c5_cycle'):
    v=d[k]; print(k, v['value'], v.get('iters_per_solve'), {kk:(vv['calls'],vv['frac']) for kk,vv in v['roofline']['kernels'].items()})
PY
