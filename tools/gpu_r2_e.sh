#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_g1_stages_gpu.py -m gpu -q -x -k "forward or build_W or coulomb or three_threads" > gpurun_out/pytest_dft.log 2>&1; echo "pytest NT=1 rc=$?"; tail -3 gpurun_out/pytest_dft.log
ISDF_DFT_NT=2 timeout 900 python -m pytest tests/test_g1_stages_gpu.py -m gpu -q -x -k "forward or build_W or coulomb or three_threads" > gpurun_out/pytest_dft2.log 2>&1; echo "pytest NT=2 rc=$?"; tail -3 gpurun_out/pytest_dft2.log
echo NT=1; timeout 300 python tools/fft_bench.py 2>&1 | tee gpurun_out/fft_bench_dft_nt1.log
echo NT=2; ISDF_DFT_NT=2 timeout 300 python tools/fft_bench.py 2>&1 | tee gpurun_out/fft_bench_dft_nt2.log
