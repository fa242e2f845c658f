#!/bin/bash
# round 2: RoeSA on the GPU (parity) + throughput of the RoeSA workload
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
run() { name=$1; shift
  timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_$name.log 2>&1; echo "bench $name rc=$?"
  tail -1 gpurun_out/bench_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], d['kernels_ms'])" || tail -5 gpurun_out/bench_$name.log
}
run roesa_n160 --workload ns3d_roesa --edge 160
