#!/bin/bash
# round 2, call A: parity at size + A/B of the pipelined Roe FD loop + ncu source-level capture + FP64 issue peak
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.max.sm,memory.total --format=csv > gpurun_out/smi.txt 2>&1
nproc > gpurun_out/nproc.txt
for n in 128 220; do
  for p in 0 1; do
    CFB200_ROE_PIPELINE=$p timeout 600 python bench.py --n $n --steps 5 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n${n}_pipe${p}.log 2>&1; echo "bench n=$n pipe=$p rc=$?"
    tail -1 gpurun_out/bench_n${n}_pipe${p}.log | python -c "import sys,json; d=json.loads(sys.stdin.read()); print(d['value']/1e6, 'Mcells/s', d['ms_per_step'], d['kernels_ms'])"
  done
done
timeout 2400 python -m pytest tests -m gpu -x -q --durations=10 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o /tmp/fp64_lat tools/microbench/fp64_lat.cu && /tmp/fp64_lat > gpurun_out/fp64_microbench.txt 2>&1; cat gpurun_out/fp64_microbench.txt
timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:InnerFaceK" -s 3 -c 1 \
    -o gpurun_out/prof_inner_n128_pipe1 -f python bench.py --n 128 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "ncu full rc=$?"
CFB200_ROE_PIPELINE=0 timeout 900 ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:InnerFaceK" -s 3 -c 1 \
    -o gpurun_out/prof_inner_n128_pipe0 -f python bench.py --n 128 --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full0.log 2>&1
echo "ncu full0 rc=$?"
ls -la gpurun_out
