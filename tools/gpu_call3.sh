#!/bin/bash
mkdir -p gpurun_out
./tools/fp64_probe > gpurun_out/fp64_probe2.txt 2>&1; tail -6 gpurun_out/fp64_probe2.txt
timeout 900 python -m pytest tests/test_stages_gpu.py -m gpu -x -q > gpurun_out/pytest_stages.log 2>&1; echo "pytest stages rc=$?"; tail -3 gpurun_out/pytest_stages.log
ISDF_GEMM_TILE=10 timeout 900 python -m pytest tests/test_stages_gpu.py -m gpu -x -q -k gemm > gpurun_out/pytest_gemm_cpasync.log 2>&1; echo "pytest cpasync rc=$?"; tail -3 gpurun_out/pytest_gemm_cpasync.log
timeout 900 python tools/gemm_bench.py 1 2 11 12 > gpurun_out/gemm_bench.txt 2>&1; echo "gemm_bench rc=$?"; cat gpurun_out/gemm_bench.txt
GEMM_CASE=flush timeout 300 python tools/gemm_bench.py 1 11 > gpurun_out/plain_gb.log 2>&1 &&
GEMM_CASE=flush timeout 900 ncu --set full --clock-control none --import-source on -k regex:dgemm_kernel -s 2 -c 1 -o gpurun_out/prof_gemm_tma python tools/gemm_bench.py 1 > gpurun_out/ncu_gb1.log 2>&1
GEMM_CASE=flush timeout 900 ncu --set full --clock-control none --import-source on -k regex:dgemm_kernel -s 2 -c 1 -o gpurun_out/prof_gemm_cpasync python tools/gemm_bench.py 11 > gpurun_out/ncu_gb2.log 2>&1
echo "ncu rc=$?"
