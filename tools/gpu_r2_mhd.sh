#!/bin/bash
# round 2: configs[3] (MHD3DProjection, q = 9) single GPU at 128^3: bench line + one --set full capture of its face kernel
mkdir -p gpurun_out
timeout 900 python bench.py --workload mhd3d --edge 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_mhd_n128.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_mhd_n128.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], d['kernels_ms'])"
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:CentredInnerFaceK" -s 3 -c 1 \
    -o gpurun_out/prof_mhd_n128 -f python bench.py --workload mhd3d --edge 128 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_mhd.log 2>&1
echo "ncu rc=$?"
