#!/bin/bash
# short check of a new binary: smoke (parity of the flagship case against the oracle) and the default bench without the CPU baseline
mkdir -p gpurun_out
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_quick.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_quick.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(round(d['value']/1e6,1), 'M cells/s', round(d['ms_per_step'],2), 'ms; e2e', round(d['e2e']['ms_per_step'],2), d['kernels_ms'])" || tail -5 gpurun_out/bench_quick.log
