#!/bin/bash
# the relinked final library (host mesh builder guard): smoke, default bench (quick), tetrahedra workload at its default edge
bash tools/gpu_r2_r.sh
timeout 600 python bench.py --workload ns3d_tet --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_ns3d_tet.log 2>&1; echo "tet rc=$?"
tail -1 gpurun_out/bench_ns3d_tet.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1), 'M cells/s', round(d['ms_per_step'],2), 'ms', d['config']['cells_per_gpu'], d['kernels_ms'])" || tail -5 gpurun_out/bench_ns3d_tet.log
