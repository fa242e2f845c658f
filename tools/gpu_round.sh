#!/bin/bash
# full evidence pass: parity tests, smoke, bench (default = configs[2]), ncu launch list, one --set full capture
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
timeout 1500 python -m pytest tests -m gpu -x -q --durations=8 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_n220.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n220.log
timeout 600 python bench.py --n 128 --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n128.log 2>&1; echo "rc=$?" >> gpurun_out/bench_n128.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_n220.csv python bench.py --steps 2 --warmup 3 > gpurun_out/ncu_launches.log 2>&1
timeout 1200 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:InnerFaceK|DiagGather|LimiterK|Gradients|TimeRhs" -s 10 -c 5 \
    -o gpurun_out/prof_full_n220 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
tail -3 gpurun_out/pytest_gpu.log gpurun_out/smoke.log gpurun_out/bench_n128.log
tail -2 gpurun_out/bench_n220.log | cut -c1-1500
timeout 600 python bench.py --workload mhd3d --n 96 --steps 5 --warmup 3 > gpurun_out/bench_mhd_n96.log 2>&1; echo "mhd rc=$?"
tail -1 gpurun_out/bench_mhd_n96.log | cut -c1-900
