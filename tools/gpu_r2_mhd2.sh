#!/bin/bash
# round 2: MHD face kernel after the reduced eigenvalue folds + batched write-out: parity (MHD cases + whole suite), bench
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
for n in 128 160; do
timeout 900 python bench.py --workload mhd3d --edge $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_mhd_n$n.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_mhd_n$n.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], d['kernels_ms'])"
done
