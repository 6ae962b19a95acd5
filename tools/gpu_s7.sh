#!/bin/bash
# GPU session 7: fused update + Gram kernel -- parity tests, kernel timing (plain vs fused), bench
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_kernels.py -m gpu -x -q -k "fused_update_gram or times_mat_add_mv or fused_update_dots" > gpurun_out/r2_pytest_ug_s7.txt 2>&1
tail -15 gpurun_out/r2_pytest_ug_s7.txt
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_s7.txt 2>&1
tail -8 gpurun_out/r2_pytest_s7.txt
out=gpurun_out/r2_kbench_ug_s7.txt; : > $out
echo "## plain kernels (B200_SPEC_GRAM=0)" >> $out
B200_SPEC_GRAM=0 timeout 300 python tools/kbench.py 1000000 4 2>&1 | grep -E "trans_mv p=(16|32|64|100|200) q=4|times_mat_add_mv p=(16|32|64|100|200) q=4" >> $out
echo "## every eligible update fused with the next Gram (B200_SPEC_GRAM=2)" >> $out
B200_SPEC_GRAM=2 timeout 300 python tools/kbench.py 1000000 4 2>&1 | grep -E "times_mat_add_mv p=(16|32|64|100|200) q=4" >> $out
echo "## q=8" >> $out
B200_SPEC_GRAM=0 timeout 300 python tools/kbench.py 1000000 8 2>&1 | grep -E "trans_mv p=(32|100|200) q=8|times_mat_add_mv p=(32|100|200) q=8" >> $out
B200_SPEC_GRAM=2 timeout 300 python tools/kbench.py 1000000 8 2>&1 | grep -E "times_mat_add_mv p=(32|100|200) q=8" >> $out
cat $out
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/r2_bench_s7.json 2> gpurun_out/r2_bench_s7.err
tail -c 1500 gpurun_out/r2_bench_s7.err
python - <<'PY'
import json
d=json.loads([l for l in open('gpurun_out/r2_bench_s7.json') if l.startswith('{')][-1])
print('value',d['value'],'e2e',d['e2e']['value'],'launches',d['gpu_launches'],'syncs',d['host_syncs'])
print({k:(v['calls'],round(v['ms'],1),v['frac']) for k,v in d['roofline']['kernels'].items()})
for k in ('north_star_8m','c5_cycle'):
    v=d[k]; print(k, v['value'], {kk:(vv['calls'],round(vv['ms'],1),vv['frac']) for kk,vv in v['roofline']['kernels'].items()})
PY
