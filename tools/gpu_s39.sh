#!/bin/bash
# session 39 (one GPU): in-solve A/B of the window SpMM producer: previous structure (build/v_old) against the
# two-descriptor-set producer with prefetch distance 1 / 2
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
out=gpurun_out/s39_spmm_insolve_ab.txt; : > $out
cp trilinos_b200/lib/libb200belos.so /tmp/lib_new.so
run() { # label workload steps
  python bench.py --steps $3 --warmup 2 --workload $2 --no-extras --no-cpu-baseline > gpurun_out/s39_tmp.json 2> gpurun_out/s39_tmp.err
  python - "$1" "$2" <<'PY' >> gpurun_out/s39_spmm_insolve_ab.txt
import json, sys
try:
    d=json.loads([l for l in open('gpurun_out/s39_tmp.json') if l.startswith('{')][-1])
    k=d['roofline']['kernels']['spmm']
    print(sys.argv[1], sys.argv[2], 'value', round(d['value'],1), 'spmm us/call', round(1000*k['ms']/k['calls'],1), 'frac', k['frac'])
except Exception as e:
    print(sys.argv[1], sys.argv[2], 'failed', e)
PY
}
cp build/v_old/trilinos_b200/lib/libb200belos.so trilinos_b200/lib/libb200belos.so
run old laplace3d_200_b4 2; run old laplace3d_100_b4 5
cp /tmp/lib_new.so trilinos_b200/lib/libb200belos.so
B200_WIN_PF=1 run new_pf1 laplace3d_200_b4 2; B200_WIN_PF=1 run new_pf1 laplace3d_100_b4 5
B200_WIN_PF=2 run new_pf2 laplace3d_200_b4 2; B200_WIN_PF=2 run new_pf2 laplace3d_100_b4 5
cp build/v_old/trilinos_b200/lib/libb200belos.so trilinos_b200/lib/libb200belos.so
run old laplace3d_200_b4 2
cp /tmp/lib_new.so trilinos_b200/lib/libb200belos.so
cat $out
