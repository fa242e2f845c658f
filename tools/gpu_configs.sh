#!/bin/bash
# bench lines of the other BASELINE configs (not the headline metric): configs[1] (2D Euler AUSM+up analytic vs FD Jacobian), configs[3] (MHD)
mkdir -p gpurun_out
for w in euler2d_analytic euler2d_numjac; do
  timeout 600 python bench.py --workload $w --n 1000 --steps 10 --warmup 3 > gpurun_out/bench_$w.log 2>&1; echo "$w rc=$?"
  tail -1 gpurun_out/bench_$w.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1),'Mcells/s', d['ms_per_step'], d['kernels_ms'])"
done
timeout 600 python bench.py --workload mhd3d --n 128 --steps 10 --warmup 3 > gpurun_out/bench_mhd3d.log 2>&1; echo "mhd rc=$?"
tail -1 gpurun_out/bench_mhd3d.log | python -c "import json,sys; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1),'Mcells/s', d['ms_per_step'], d['kernels_ms'])"
