#!/bin/bash
# parity tests + the Corrected3D workload + default bench
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -x -q --durations=5 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -4 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload ns3d_corrected --steps 10 --warmup 3 > gpurun_out/bench_ns3d_corrected.log 2>&1; echo "rc=$?"
tail -1 gpurun_out/bench_ns3d_corrected.log | cut -c1-1800
timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n220b.log 2>&1; echo "rc=$?"
tail -1 gpurun_out/bench_n220b.log | cut -c1-300
