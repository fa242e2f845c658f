#!/bin/bash
# round 2, call B: parity tests + fused time term + stored LSQ weights + epilogue batching, bench at 128 / 220
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q --durations=6 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -12 gpurun_out/pytest_gpu.log
for n in 128 220; do
  timeout 600 python bench.py --n $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n${n}.log 2>&1; echo "bench n=$n rc=$?"
  tail -1 gpurun_out/bench_n${n}.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['kernels_ms'])" || tail -5 gpurun_out/bench_n${n}.log
done
timeout 600 python bench.py --n 220 --steps 5 --warmup 3 --no-cpu-baseline --no-fuse > gpurun_out/bench_n220_nofuse.log 2>&1
tail -1 gpurun_out/bench_n220_nofuse.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('nofuse', d['value']/1e6, 'Mcells/s', d['ms_per_step'], d['kernels_ms'])"
timeout 900 python bench.py --n 220 --steps 3 --warmup 1 > gpurun_out/bench_n220_cpu.log 2>&1; echo "bench cpu rc=$?"
tail -1 gpurun_out/bench_n220_cpu.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['cpu_baseline'])"
(time timeout 900 python bench.py --impl reference --n 220 --steps 3 --warmup 1) > gpurun_out/bench_ref.log 2>&1; tail -5 gpurun_out/bench_ref.log | cut -c1-600
