#!/bin/bash
# Round 2, call S: the TMA-fed GEMM as default -- full GPU suite, contract bench at N = 1, GEMM shapes against the
# LDGSTS kernel and cuBLAS, launch list of one C64 step, ncu --set full of the flush on the new kernel.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=6 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/pytest_gpu.log
timeout 200 python tools/gemm_tma_check.py > gpurun_out/gemm_tma_check.log 2>&1; echo "tma check rc=$?"; tail -2 gpurun_out/gemm_tma_check.log
timeout 300 python tools/gemm_bench.py 4 0 > gpurun_out/gemm_bench_tma.log 2>&1; cat gpurun_out/gemm_bench_tma.log
timeout 900 python bench.py --steps 5 --warmup 3 > gpurun_out/bench_n1_tma.json 2> gpurun_out/bench_n1_tma.err; echo "bench rc=$?"; tail -c 2500 gpurun_out/bench_n1_tma.json
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches_c64_tma.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
python tools/summarize_launches.py gpurun_out/launches_c64_tma.csv gpurun_out/launches_c64_tma.md | head -30
GEMM_CASE=flush timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_tma_kernel -c 2 -f -o gpurun_out/prof_dgemm_tma_r02 python tools/gemm_bench.py 0 > gpurun_out/ncu_dgemm_tma.log 2>&1; echo "ncu gemm rc=$?"; tail -3 gpurun_out/ncu_dgemm_tma.log
