#!/bin/bash
# round 2 evidence pass on the shipped binary: smoke, default bench line (+ cpu_baseline), reference arm, ncu launch list,
# one --set full capture of the heaviest kernels of the default bench command
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
timeout 300 python __graft_entry__.py smoke > gpurun_out/smoke.log 2>&1; echo "smoke rc=$?" >> gpurun_out/smoke.log
tail -2 gpurun_out/smoke.log
timeout 900 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_default.log 2>&1; echo "bench rc=$?"
tail -1 gpurun_out/bench_default.log | cut -c1-200
timeout 900 python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/bench_reference.log 2>&1; echo "reference rc=$?"
tail -1 gpurun_out/bench_reference.log | cut -c1-300
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_n220.csv python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_launches.log 2>&1
echo "ncu launches rc=$?"
timeout 1500 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:InnerFaceK|DiagGather|LimiterK|Gradients|CellGather" -s 10 -c 6 \
    -o gpurun_out/prof_full_n220 -f python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
ls -la gpurun_out/*.ncu-rep
