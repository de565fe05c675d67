#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fft_ -c 3 -f -o gpurun_out/prof_fft python tools/fft_bench.py 71 > gpurun_out/ncu_fft.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_fft.log
