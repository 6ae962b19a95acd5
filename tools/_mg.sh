N=$1
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
python -m pytest tests/test_multigpu_nccl.py -m gpu -x -q > gpurun_out/s23_test_n$N.log 2>&1; tail -3 gpurun_out/s23_test_n$N.log
$TR --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 > gpurun_out/s23_n${N}_c2.json 2> gpurun_out/s23_n${N}_c2.err
$TR --master-port 29513 bench.py --gpus $N --steps 2 --warmup 3 --workload laplace3d_200_b4 > gpurun_out/s23_n${N}_c3.json 2> gpurun_out/s23_n${N}_c3.err
