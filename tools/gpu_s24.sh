#!/bin/bash
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2_pytest_s24.txt 2>&1
tail -25 gpurun_out/r2_pytest_s24.txt
timeout 800 python bench.py --workload brick3d_150_b8 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/r2_bench_c4.json 2> gpurun_out/r2_bench_c4.err; tail -c 300 gpurun_out/r2_bench_c4.err
python - <<PY
import json
d=json.loads([l for l in open("gpurun_out/r2_bench_c4.json") if l.startswith("{")][-1])
print("value",d["value"],"e2e",d["e2e"]["value"],"iters",d["run_info"]["iters_per_solve"],"launches",d["gpu_launches"],"syncs",d["host_syncs"])
print({k:(v["calls"],round(v["ms"],1),v["frac"]) for k,v in d["roofline"]["kernels"].items()})
print(d["roofline"]["solve"])
PY
