#!/bin/bash
# The driver's GPU test command, then the no-allocation entry point alone in a fresh process.
mkdir -p gpurun_out
timeout 1800 python -m pytest tests -m gpu -q -x --durations=10 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -22 gpurun_out/pytest_gpu.log
timeout 600 python -m pytest tests/test_zzz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/pytest_noalloc_alone.log
