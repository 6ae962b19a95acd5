# usage: bash tools/scale_run.sh N TAG   (run under gpurun --gpus N): default workload and the 8M-row problem on N GPUs
N=$1; TAG=$2
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1"
$TR --master-port 29511 bench.py --gpus $N --steps 3 --warmup 3 2> gpurun_out/${TAG}_n${N}_c2.err | grep '^{' > gpurun_out/${TAG}_n${N}_c2.json
$TR --master-port 29513 bench.py --gpus $N --steps 2 --warmup 3 --workload laplace3d_200_b4 2> gpurun_out/${TAG}_n${N}_c3.err | grep '^{' > gpurun_out/${TAG}_n${N}_c3.json
