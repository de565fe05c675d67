#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -15 gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload c64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c64.json 2> gpurun_out/bench_c64.err; echo "c64 rc=$?"; tail -3 gpurun_out/bench_c64.err
python -c "
import json; d=json.loads(open('gpurun_out/bench_c64.json').read().strip().splitlines()[-1]); print(d['value'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d['roofline']['achieved'], d['roofline']['gemm_share_of_step'])"
for kb in 96 48 32 24; do echo "FFT smem budget $kb KB"; ISDF_FFT_SMEM_KB=$kb timeout 300 python tools/fft_bench.py; done > gpurun_out/fft_bench.txt 2>&1
cat gpurun_out/fft_bench.txt
