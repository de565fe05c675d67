#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_kernel -c 2 -f -o gpurun_out/prof_dgemm_final python tools/gemm_bench.py 0 > gpurun_out/ncu_dgemm.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_dgemm.log
