#!/bin/bash
# round 2, call D: L2 prefetch of the next wave's gathers (CFB200_PREFETCH_AHEAD), separate vs fused time term
mkdir -p gpurun_out
for pf in 0 37888 75776; do
  CFB200_PREFETCH_AHEAD=$pf timeout 600 python bench.py --n 220 --steps 5 --warmup 3 --no-cpu-baseline --no-fuse > gpurun_out/bench_n220_pf${pf}.log 2>&1; echo "bench pf=$pf rc=$?"
  tail -1 gpurun_out/bench_n220_pf${pf}.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['kernels_ms'])" || tail -5 gpurun_out/bench_n220_pf${pf}.log
done
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
