#!/bin/bash
# session 35: the allocation-free entry point alone in a fresh process, then the whole GPU suite (it runs last), then the bench
mkdir -p gpurun_out
python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/s35_noalloc_alone.txt 2>&1; echo "alone rc=$?" | tee -a gpurun_out/s35_noalloc_alone.txt
python -m pytest tests -m gpu -x -q > gpurun_out/s35_pytest_gpu.txt 2>&1; echo "suite rc=$?" | tee -a gpurun_out/s35_pytest_gpu.txt
tail -3 gpurun_out/s35_noalloc_alone.txt gpurun_out/s35_pytest_gpu.txt
python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/s35_smoke.txt 2>&1; echo "smoke rc=$?"
python bench.py > gpurun_out/s35_bench_n1.json 2> gpurun_out/s35_bench_n1.err; echo "bench rc=$?"; cat gpurun_out/s35_bench_n1.json
