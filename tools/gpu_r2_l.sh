#!/bin/bash
# round 2: ncu --set full of the analytic face kernel, configs[1] at 2048^2
mkdir -p gpurun_out
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:AnalyticInnerFaceK" -s 3 -c 1 \
    -o gpurun_out/prof_e2d_analytic -f python bench.py --workload euler2d_analytic --edge 2048 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_e2d.log 2>&1
echo "ncu rc=$?"
