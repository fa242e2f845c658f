#!/bin/bash
# parity tests + bench at 128^3 and at the default (configs[2]) size
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -n 3 gpurun_out/pytest_gpu.log
for n in ${SIZES:-128 220}; do
  timeout 900 python bench.py --n $n --steps ${STEPS:-10} --warmup 3 --no-cpu-baseline > gpurun_out/bench_n$n.log 2>&1
  python - <<PY
import json
try:
    d=json.loads(open('gpurun_out/bench_n$n.log').read().strip().splitlines()[-1])
    print('n $n', round(d['value']/1e6,1),'Mcells/s', round(d['ms_per_step'],3), 'e2e', round(d['e2e']['value']/1e6,1), d['kernels_ms'])
except Exception as e:
    print('n $n failed', e)
PY
done
