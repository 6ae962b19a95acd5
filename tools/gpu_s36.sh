#!/bin/bash
# session 36, N = $1 ranks: multi-rank parity, then A/B of the merged apply with the descriptors in list order (default)
# against the index list (B200_SPMM_MIX_LIST=1) on the 8M-row problem, then config 2
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_multigpu_nccl.py -m gpu -x -q > gpurun_out/s36_pytest_multigpu_n$N.txt 2>&1; echo "multigpu rc=$?"
tail -n 2 gpurun_out/s36_pytest_multigpu_n$N.txt
i=0
for v in "B200_X=0" "B200_SPMM_MIX_LIST=1"; do
  for w in laplace3d_200_b4 laplace3d_100_b4; do
  i=$((i+1))
  env $v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2962$i \
    bench.py --gpus $N --steps 2 --warmup 3 --workload $w --no-extras --no-cpu-baseline > gpurun_out/s36_ab_${i}_n$N.json 2> gpurun_out/s36_ab_${i}_n$N.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/s36_ab_${i}_n$N.json') if l.startswith('{')][-1])
    k=d['roofline']['kernels']
    print('$v', '$w', 'value', round(d['value'],1), 'spmm calls', k['spmm']['calls'], 'us/call', round(1000*k['spmm']['ms']/k['spmm']['calls'],1), 'frac', k['spmm']['frac'])
except Exception as e:
    print('$v $w failed', e)
PY
  done
done | tee gpurun_out/s36_apply_ab_n$N.txt
