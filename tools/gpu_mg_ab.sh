#!/bin/bash
# multi-GPU A/B of the operator apply on the 8M-row problem; N = $1 ranks
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
i=0
for v in "B200_X=0" "B200_NO_FUSE=1" "B200_NO_P2P=1" "B200_SPMM_WIN=0"; do
  i=$((i+1))
  env $v timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2952$i \
    bench.py --gpus $N --steps 1 --warmup 3 --workload laplace3d_200_b4 --no-extras --no-cpu-baseline > gpurun_out/r2_ab_${i}_n$N.json 2> gpurun_out/r2_ab_${i}_n$N.err
  python - <<PY
import json
try:
    d=json.loads([l for l in open('gpurun_out/r2_ab_${i}_n$N.json') if l.startswith('{')][-1])
    k=d['roofline']['kernels']
    print('$v', 'value', round(d['value'],1), 'spmm', k['spmm']['calls'], round(k['spmm']['ms'],1), k['spmm']['frac'], 'halo', k.get('halo'))
except Exception as e:
    print('$v failed', e)
PY
done
