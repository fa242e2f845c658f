#!/bin/bash
# round 2: RHS-only (explicit) step with the face kernel at three resident blocks per SM; analytic workloads re-measured
mkdir -p gpurun_out
run() { name=$1; shift
  timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_$name.log 2>&1; echo "bench $name rc=$?"
  tail -1 gpurun_out/bench_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], 'step frac', d['step_roofline']['frac_of_hbm_peak'], d['kernels_ms'])" || tail -5 gpurun_out/bench_$name.log
}
run ns3d_rhs_only_n220 --edge 220 --rhs-only
run ns3d_rhs_only_n317 --edge 317 --rhs-only
run euler2d_analytic_n4096 --workload euler2d_analytic --edge 4096
run ausm_analytic_n220 --workload ns3d_ausm_analytic --edge 220
timeout 1200 python -m pytest tests/test_gpu_parity.py -m gpu -x -q > gpurun_out/pytest_gpu_parity.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu_parity.log
tail -3 gpurun_out/pytest_gpu_parity.log
