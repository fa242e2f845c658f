#!/bin/bash
# N GPUs of one box: NCCL tests, weak scaling (configs[2], configs[3] MHD), one strong-scaling point
N=${1:-8}
mkdir -p gpurun_out
nvidia-smi -L > gpurun_out/smi_${N}gpu.txt
timeout 900 python -m pytest tests/test_gpu_multi.py -x -q > gpurun_out/pytest_gpu_multi_${N}gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_multi_${N}gpu.log
tail -4 gpurun_out/pytest_gpu_multi_${N}gpu.log
run() { name=$1; shift
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $N --master-addr 127.0.0.1 --master-port 29517 bench.py --gpus $N --steps 5 --warmup 3 "$@" > gpurun_out/bench_${N}gpu_$name.log 2>&1; rc=$?
  grep '^{"metric"' gpurun_out/bench_${N}gpu_$name.log | tail -1 | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', round(d['value']/1e6,1), 'Mcells/s', round(d['ms_per_step'],3), 'ms e2e', round(d['e2e']['ms_per_step'],2), d['scaling'], (d.get('parity') or {}).get('ok'))" || { echo "$name rc=$rc"; tail -15 gpurun_out/bench_${N}gpu_$name.log; }
}
run weak_ns3d_n220 --edge 220
run weak_mhd3d_n128 --edge 128 --workload mhd3d --no-parity
run strong_ns3d_n224 --edge 224 --strong --no-parity
