#!/bin/bash
# N-GPU pass: NCCL halo tests + weak-scaling bench line at N GPUs. usage: gpu_multi.sh N [n]
N=${1:-2}; n=${2:-128}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_multi.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -m gpu -x -q > gpurun_out/pytest_multi.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_multi.log
tail -n 15 gpurun_out/pytest_multi.log
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus $N --edge $n --steps 10 --warmup 3 > gpurun_out/bench_g${N}_n$n.log 2>&1; echo "rc=$?" >> gpurun_out/bench_g${N}_n$n.log
tail -n 3 gpurun_out/bench_g${N}_n$n.log | cut -c1-600
