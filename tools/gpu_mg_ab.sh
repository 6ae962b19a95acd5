#!/bin/bash
# multi-GPU A/B: one SpMM launch (interior + boundary tiles) vs two; N = $1 ranks
N=${1:-2}
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_multigpu_nccl.py -m gpu -x -q -k "$N" > gpurun_out/r2_pytest_mg_ab_n$N.txt 2>&1
tail -3 gpurun_out/r2_pytest_mg_ab_n$N.txt
for split in 0 1; do
  B200_SPMM_SPLIT_LAUNCH=$split timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 2951$split \
    bench.py --gpus $N --steps 2 --warmup 3 --workload laplace3d_200_b4 --no-extras --no-cpu-baseline > gpurun_out/r2_ab_split${split}_n$N.json 2> gpurun_out/r2_ab_split${split}_n$N.err
  python - <<PY
import json
d=json.loads([l for l in open('gpurun_out/r2_ab_split${split}_n$N.json') if l.startswith('{')][-1])
k=d['roofline']['kernels']
print('split_launch=${split}', 'value', round(d['value'],1), 'spmm', k['spmm'])
PY
done
