#!/bin/bash
# N GPUs of one box: NCCL halo tests (bitwise vs single domain, Jacobian rows included) + bench with the cross-rank parity field
N=${1:-2}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_multi.txt
timeout 1800 python -m pytest tests/test_gpu_multi.py tests/test_gpu_solver.py -x -q --durations=5 > gpurun_out/pytest_gpu_multi_${N}gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_multi_${N}gpu.log
tail -12 gpurun_out/pytest_gpu_multi_${N}gpu.log
for n in ${SIZES:-128 220}; do
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --edge $n --steps 5 --warmup 3 > gpurun_out/bench_${N}gpu_n${n}.log 2>&1; echo "bench N=$N n=$n rc=$?"
  tail -1 gpurun_out/bench_${N}gpu_n${n}.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d.get('parity'))" || tail -20 gpurun_out/bench_${N}gpu_n${n}.log
done
