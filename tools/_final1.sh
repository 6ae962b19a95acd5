mkdir -p gpurun_out
python -m pytest tests -m gpu -x -q > gpurun_out/s24_pytest.log 2>&1; echo "rc=$?" >> gpurun_out/s24_pytest.log
python bench.py --steps 5 --warmup 3 > gpurun_out/s24_bench.json 2> gpurun_out/s24_bench.err
python bench.py --steps 2 --warmup 3 --workload laplace3d_200_b4 --no-cpu-baseline > gpurun_out/s24_bench_c3.json 2> gpurun_out/s24_bench_c3.err
python tools/kbench.py > gpurun_out/s24_kbench.txt 2>&1
python tools/profile_cycle.py laplace3d_100_b4 1 > gpurun_out/s24_cycle.json 2>&1 && ncu --metrics gpu__time_duration.sum --clock-control none -c 6000 --csv --log-file gpurun_out/s24_launches.csv python tools/profile_cycle.py laplace3d_100_b4 1 > gpurun_out/s24_ncu_cycle.log 2>&1
python tools/ncu_kernels.py > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -o gpurun_out/s24_kernels python tools/ncu_kernels.py > gpurun_out/s24_ncu.log 2>&1
ncu -i gpurun_out/s24_kernels.ncu-rep --page raw --csv > gpurun_out/s24_kernels_raw.csv 2>/dev/null
python tools/ncu_summary.py gpurun_out/s24_kernels.ncu-rep > gpurun_out/s24_kernels_summary.txt 2>&1
