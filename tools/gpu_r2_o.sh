#!/bin/bash
# round 2: two-part evaluation with overlapped host copies (default) against one part (CFB200_SPLIT_D2H=0): e2e of the default bench
mkdir -p gpurun_out
run() { name=$1; shift
  timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline "$@" > gpurun_out/bench_$name.log 2>&1; echo "bench $name rc=$?"
  tail -1 gpurun_out/bench_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1), 'Mcells/s', round(d['ms_per_step'],2), 'e2e ms', round(d['e2e']['ms_per_step'],2), 'e2e M', round(d['e2e']['value']/1e6,1))" || tail -5 gpurun_out/bench_$name.log
}
for s in 1 0 1 0; do
  export CFB200_SPLIT_D2H=$s
  run ns3d_n220_split$s --edge 220
done
export CFB200_SPLIT_D2H=1
run mhd_n128_split1 --workload mhd3d --edge 128
run e2d_split1 --workload euler2d_analytic --edge 4096
export CFB200_SPLIT_D2H=0
run mhd_n128_split0 --workload mhd3d --edge 128
run e2d_split0 --workload euler2d_analytic --edge 4096
unset CFB200_SPLIT_D2H
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
