#!/bin/bash
# tests + FFT microbench + default bench (N=1) + 2-GPU c64 with the stage breakdown
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
timeout 300 python tools/fft_bench.py > gpurun_out/fft_bench2.txt 2>&1; cat gpurun_out/fft_bench2.txt
export ISDF_PROFILE_SELECT=1
timeout 600 python bench.py --workload c64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c64.json 2> gpurun_out/bench_c64.err; echo "c64 rc=$?"; tail -3 gpurun_out/bench_c64.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c64.json').read().strip().splitlines()[-1]); print(d['value'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d['roofline']['achieved'], d['roofline']['gemm_share_of_step'])"
