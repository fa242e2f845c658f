#!/bin/bash
mkdir -p gpurun_out
run() { name=$1; shift
  timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_$name.log 2>&1; echo "bench $name rc=$?"
  tail -1 gpurun_out/bench_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], d['kernels_ms'])" || tail -5 gpurun_out/bench_$name.log
}
run ns3d_n220 --edge 220
run tet_n100 --workload ns3d_tet --edge 100
