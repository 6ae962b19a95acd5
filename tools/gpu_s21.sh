#!/bin/bash
# GPU session 21: MPI probe on the GPU box, config-5 progress over 15 restart cycles, ncu launch list of one config-2
# cycle, ncu --set full of the fused update + Gram kernel and of the findBasis column step
cd "$(dirname "$0")/.."
mkdir -p gpurun_out
{ echo "# which mpirun mpiexec mpicc orterun srun (GPU box)"; which mpirun mpiexec mpicc orterun srun; echo "exit code $?"; ls /usr/lib/x86_64-linux-gnu 2>/dev/null | grep -i -E "mpi|openmpi|mpich" ; echo "(end of library listing)"; nproc; } > gpurun_out/r2_mpirun_probe.txt 2>&1
cat gpurun_out/r2_mpirun_probe.txt
timeout 300 python tools/c5_progress.py 15 > gpurun_out/r2_c5_progress_s21.txt 2>&1
cat gpurun_out/r2_c5_progress_s21.txt | tail -22
python tools/profile_cycle.py laplace3d_100_b4 1 > gpurun_out/r2_cycle_plain.log 2>&1 && \
ncu --metrics gpu__time_duration.sum --clock-control none --csv --log-file gpurun_out/r2_launches_config2_cycle.csv python tools/profile_cycle.py laplace3d_100_b4 1 > gpurun_out/r2_cycle_ncu.log 2>&1
tail -2 gpurun_out/r2_cycle_plain.log
B200_SPEC_GRAM=2 python tools/ncu_ug.py 1000000 100 4 2 > gpurun_out/ncu_ug_plain.log 2>&1 && \
B200_SPEC_GRAM=2 ncu --set full --clock-control none --import-source on -k regex:update_gram -s 3 -c 1 -o gpurun_out/r2_ug100_s21 -f python tools/ncu_ug.py 1000000 100 4 2 > gpurun_out/ncu_ug100.log 2>&1
tail -1 gpurun_out/ncu_ug100.log
python tools/ncu_kernels.py > gpurun_out/ncu_k_plain.log 2>&1 && \
ncu --set full --clock-control none -k regex:"fb_step|gram_tma|update_kernel|spmm_win" -s 8 -c 8 -o gpurun_out/r2_kernels_s21 -f python tools/ncu_kernels.py > gpurun_out/ncu_k.log 2>&1
tail -1 gpurun_out/ncu_k.log
