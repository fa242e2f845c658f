#!/bin/bash
# parity tests + launch-shape sweep of the perturbation kernel
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
for s in ${SHAPES:-2 3 4 5}; do
  CFB200_FACE_SHAPE=$s timeout 600 python bench.py --n ${N:-128} --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_shape$s.log 2>&1
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_shape$s.log').read().strip().splitlines()[-1])
    print('shape $s', round(d['value']/1e6,1),'Mcells/s', d['ms_per_step'], d['kernels_ms'])
except Exception as e:
    print('shape $s failed', e)
PY
done
