#!/bin/bash
# ncu --set full captures of the three FFT passes (default and fused kernels) on the 71^3 mesh, after the
# same command has exited 0 without ncu; then the microbench numbers for every BASELINE mesh, both forms.
mkdir -p gpurun_out
timeout 300 python tools/fft_bench.py > gpurun_out/fft_bench_default.log 2>&1; echo "fft bench rc=$?"; cat gpurun_out/fft_bench_default.log
ISDF_FFT_FUSED=1 timeout 300 python tools/fft_bench.py > gpurun_out/fft_bench_fused.log 2>&1; echo "fft bench fused rc=$?"; cat gpurun_out/fft_bench_fused.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fft_ -c 3 -f -o gpurun_out/prof_fft_default \
    python tools/fft_bench.py 71 > gpurun_out/ncu_fft_default.log 2>&1; echo "ncu default rc=$?"; tail -2 gpurun_out/ncu_fft_default.log
ISDF_FFT_FUSED=1 timeout 600 ncu --set full --clock-control none --import-source on -k regex:fft_ -c 3 -f -o gpurun_out/prof_fft_fused \
    python tools/fft_bench.py 71 > gpurun_out/ncu_fft_fused.log 2>&1; echo "ncu fused rc=$?"; tail -2 gpurun_out/ncu_fft_fused.log
