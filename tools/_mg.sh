python -m pytest tests/test_multigpu_nccl.py -m gpu -x -q 2>&1 | tail -5
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/s21_n2_c2.json 2> gpurun_out/s21_n2_c2.err
$TR --master-port 29513 bench.py --gpus 2 --steps 2 --warmup 3 --workload laplace3d_200_b4 > gpurun_out/s21_n2_c3.json 2> gpurun_out/s21_n2_c3.err
tail -3 gpurun_out/s21_n2_c2.err
