#!/bin/bash
# round 2: allocation-free norm entry point alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -x -q > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "alone rc=$?" >> gpurun_out/pytest_noalloc_alone.log
tail -4 gpurun_out/pytest_noalloc_alone.log
timeout 1800 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -5 gpurun_out/pytest_gpu.log
