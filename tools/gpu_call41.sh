#!/bin/bash
mkdir -p gpurun_out
timeout 600 python tools/gemm_bench.py 0 4 5 6 > gpurun_out/gemm_bench_mi4b.txt 2>&1; cat gpurun_out/gemm_bench_mi4b.txt
