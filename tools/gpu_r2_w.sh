#!/bin/bash
# ncu --set full of the ANALYTIC face kernel (configs[2] with NumJacobAnalytic at 220^3): where the workload without a
# finite-difference loop stands against DRAM and the FP64 pipe.  Only the raw-page CSV travels back (the report is > 40 MB).
mkdir -p gpurun_out
timeout 300 ncu --set full --clock-control none --kernel-name-base demangled -k "regex:AnalyticInnerFaceK" -s 3 -c 1 \
    -o /tmp/prof_analytic_ns3d -f python bench.py --workload ns3d_ausm_analytic --steps 1 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_analytic_ns3d.log 2>&1
echo "ncu ns3d analytic rc=$?"
ncu -i /tmp/prof_analytic_ns3d.ncu-rep --page raw --csv > gpurun_out/raw_analytic_ns3d.csv 2>/dev/null; ls -la gpurun_out/raw_analytic_ns3d.csv
