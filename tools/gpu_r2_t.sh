#!/bin/bash
# final binary, the other workloads (no CPU baseline): configs[1] analytic at 4096^2, configs[3] MHD at 128^3, configs[2] with the
# analytic Jacobians at 220^3, RHS only at 220^3, tetrahedra
mkdir -p gpurun_out
run() { name=$1; shift
  timeout 600 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_$name.log 2>&1; echo "bench $name rc=$?"
  tail -1 gpurun_out/bench_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1), 'Mcells/s', round(d['ms_per_step'],2), 'ms; e2e', round(d['e2e']['ms_per_step'],2), d['kernels_ms'])" || tail -5 gpurun_out/bench_$name.log
}
run euler2d_analytic_n4096 --workload euler2d_analytic --edge 4096
run mhd3d_n128 --workload mhd3d --edge 128
run ausm_analytic_n220 --workload ns3d_ausm_analytic --edge 220
run ns3d_rhs_only_n220 --rhs-only --edge 220
run ns3d_tet --workload ns3d_tet
