#!/bin/bash
# C5 sweep: hex boxes, RHS+Jacobian and RHS-only, one GPU; tet line; configs[1] analytic at size; MHD
mkdir -p gpurun_out
run() { name=$1; shift
  timeout 1200 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/sweep_$name.log 2>&1; rc=$?
  tail -1 gpurun_out/sweep_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print('$name', round(d['value']/1e6,1), 'Mcells/s', round(d['ms_per_step'],3), 'ms  step-roofline', round(d['step_roofline']['frac_of_hbm_peak'],4), d['kernels_ms'])" || { echo "$name rc=$rc"; tail -3 gpurun_out/sweep_$name.log; }
}
for n in 100 126 159 200 252 317; do
  run jac_n$n --n $n --no-fuse
  run rhs_n$n --n $n --rhs-only
done
run tet_n100 --workload ns3d_tet --n 100 --no-fuse
run tet_rhs_n100 --workload ns3d_tet --n 100 --rhs-only
run euler2d_analytic_n4096 --workload euler2d_analytic --n 4096 --no-fuse
run euler2d_numjac_n4096 --workload euler2d_numjac --n 4096 --no-fuse
run mhd3d_n128 --workload mhd3d --n 128 --no-fuse
run corrected_n220 --workload ns3d_corrected --n 220 --no-fuse
