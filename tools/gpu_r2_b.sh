#!/bin/bash
# Round 2, call B: BASELINE-size parity test file, host facts, C64 parity tool with a 2048-pivot oracle prefix.
mkdir -p gpurun_out
nproc; free -g | head -2; grep -m1 "model name" /proc/cpuinfo; grep -c avx512f /proc/cpuinfo
timeout 900 python -m pytest tests/test_g4_baseline_parity_gpu.py -m gpu -q -x --durations=8 > gpurun_out/pytest_g4.log 2>&1; echo "g4 rc=$?"; tail -25 gpurun_out/pytest_g4.log
timeout 1200 python tools/c64_parity.py --oracle-points 2048 > gpurun_out/c64_parity_2048.json 2> gpurun_out/c64_parity_2048.err; echo "c64 parity rc=$?"; tail -1 gpurun_out/c64_parity_2048.json; tail -5 gpurun_out/c64_parity_2048.err
