#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/r2_spec_threshold_s30.txt; : > $out
for nx in 40 50 63 79; do for sp in 0 1; do
  echo "## nx=$nx B200_SPEC_GRAM=$sp" >> $out
  B200_SPEC_GRAM=$sp python tools/solve_demo.py Laplace3D $nx 4 50 DGKS 1e-8 - 3 2>&1 | grep iters_per_s | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['num_iters'], round(d['iters_per_s'],1))" >> $out
done; done
cat $out
