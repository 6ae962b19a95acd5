#!/bin/bash
# 2-GPU bisect: which of the round-2 fusions desynchronises the exchange on the 8M-row problem?
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
run() {
  tag=$1; shift
  env "$@" timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 \
    bench.py --gpus 2 --steps 1 --warmup 1 --workload laplace3d_200_b4 --no-extras --no-cpu-baseline > gpurun_out/bisect_$tag.json 2> gpurun_out/bisect_$tag.err
  echo "== $tag rc=$? $(grep -c 'did not arrive' gpurun_out/bisect_$tag.err) $(head -c 200 gpurun_out/bisect_$tag.json)"
  grep -m3 "returned\|Error" gpurun_out/bisect_$tag.err
}
run both_on B200_SPEC_GRAM=1 B200_FB_STEP=1
run both_on_verbose B200_SPEC_GRAM=1 B200_FB_STEP=1 B200_VERBOSE=1
