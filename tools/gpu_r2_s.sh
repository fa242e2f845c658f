#!/bin/bash
# two GPUs of one box, final binary: NCCL halo tests (bitwise vs the single domain, Jacobian rows included) and the weak-scaling
# bench line with its cross-rank parity field
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_multi.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q --durations=5 > gpurun_out/pytest_gpu_multi_2gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_multi_2gpu.log
tail -6 gpurun_out/pytest_gpu_multi_2gpu.log
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus 2 --steps 5 --warmup 3 > gpurun_out/bench_2gpu_n220.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_2gpu_n220.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d.get('parity'))" || tail -20 gpurun_out/bench_2gpu_n220.log
