#!/bin/bash
mkdir -p gpurun_out
run() { # name, env..., args
  name=$1; shift
  env "$@" timeout 600 python bench.py --n 220 --steps 5 --warmup 3 --no-cpu-baseline $EXTRA > gpurun_out/bench_$name.log 2>&1; echo "bench $name rc=$?"
  tail -1 gpurun_out/bench_$name.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['kernels_ms'])" || tail -5 gpurun_out/bench_$name.log
}
EXTRA="--no-fuse" run nofuse_lim0 CFB200_LIMITER_SHAPE=0
EXTRA="--no-fuse" run nofuse_lim1 CFB200_LIMITER_SHAPE=1
EXTRA="" run fuse_lim0 CFB200_LIMITER_SHAPE=0
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
