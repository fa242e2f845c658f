#!/bin/bash
# round 2, call C: two-phase 3-blocks/SM Roe kernel (CFB200_ROE_PIPELINE=2) against the 2-block pipelined one (=1)
mkdir -p gpurun_out
for n in 128 220; do
  for p in 1 2; do
    CFB200_ROE_PIPELINE=$p timeout 600 python bench.py --n $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n${n}_pipe${p}.log 2>&1; echo "bench n=$n pipe=$p rc=$?"
    tail -1 gpurun_out/bench_n${n}_pipe${p}.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], 'e2e', d['e2e']['ms_per_step'], d['kernels_ms'])" || tail -5 gpurun_out/bench_n${n}_pipe${p}.log
  done
done
timeout 1800 python -m pytest tests -m gpu -x -q --durations=4 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -8 gpurun_out/pytest_gpu.log
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:InnerFacePhased" -s 3 -c 1 \
    -o gpurun_out/prof_inner_n128_pipe2 -f python bench.py --n 128 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full2.log 2>&1
echo "ncu rc=$?"
