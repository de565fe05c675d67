#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/fft_bench4.txt
for z in 24 32 48 64 96; do for c in 64 96 128; do
echo "== z=$z KB cols=$c KB" >> gpurun_out/fft_bench4.txt
ISDF_FFT_SMEM_Z_KB=$z ISDF_FFT_SMEM_KB=$c timeout 100 python tools/fft_bench.py 71 108 >> gpurun_out/fft_bench4.txt 2>&1
done; done
cat gpurun_out/fft_bench4.txt
