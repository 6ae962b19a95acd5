python tools/spmm_bench.py Laplace3D 100 4 cold > /dev/null 2>&1
ncu --set full --clock-control none --import-source on -k regex:spmm_stream -s 5 -c 1 -o gpurun_out/s18_spmm python tools/spmm_bench.py Laplace3D 100 4 cold > gpurun_out/s18_ncu.log 2>&1
ncu -i gpurun_out/s18_spmm.ncu-rep --page source --csv > gpurun_out/s18_spmm_source.csv 2>/dev/null
ncu -i gpurun_out/s18_spmm.ncu-rep --page details > gpurun_out/s18_spmm_details.txt 2>/dev/null
