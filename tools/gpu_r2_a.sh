#!/bin/bash
# Round 2, call A: the no-allocation entry point alone in a fresh process, the whole GPU suite (-x, driver's
# command), then the FFT microbench and the ncu --set full captures of the three FFT passes on the 71^3 mesh.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zzz_no_alloc.py -m gpu -q -x > gpurun_out/pytest_noalloc_alone.log 2>&1; echo "noalloc alone rc=$?"; tail -3 gpurun_out/pytest_noalloc_alone.log
timeout 1500 python -m pytest tests -m gpu -q -x --durations=15 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -25 gpurun_out/pytest_gpu.log
timeout 300 python tools/fft_bench.py > gpurun_out/fft_bench_default.log 2>&1; echo "fft bench rc=$?"; cat gpurun_out/fft_bench_default.log
timeout 600 ncu --set full --clock-control none --import-source on -k regex:fft_ -c 3 -f -o gpurun_out/prof_fft_default \
    python tools/fft_bench.py 71 > gpurun_out/ncu_fft_default.log 2>&1; echo "ncu default rc=$?"; tail -2 gpurun_out/ncu_fft_default.log
