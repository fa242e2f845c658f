#!/bin/bash
# round 2: full GPU suite after CorrectedGG / 2DNoLimiterID + default bench sanity
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_default.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_default.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], d['kernels_ms'])"
