#!/bin/bash
# one --set full capture of selected kernels at N^3: usage gpu_ncu_full.sh N 'regex' skip count
mkdir -p gpurun_out
N=${1:-128}; RX=${2:-InnerPertK|InnerFaceK}; SKIP=${3:-0}; CNT=${4:-2}
python bench.py --n $N --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
ncu --set full --clock-control none --import-source on --kernel-name-base demangled -k "regex:$RX" -s $SKIP -c $CNT \
    -o gpurun_out/prof_full -f python bench.py --n $N --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_full.log 2>&1
echo "rc=$?"; tail -n 2 gpurun_out/plain.log | cut -c1-300
