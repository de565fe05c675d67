#!/bin/bash
# Round 2, call D: dense-DFT forward passes: tests, microbench per mesh (DFT vs Stockham), C64 bench line, ncu.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_g1_stages_gpu.py tests/test_g3_isdf_gpu.py tests/test_zzy_dist_gpu.py -m gpu -q -x > gpurun_out/pytest_dft.log 2>&1; echo "pytest rc=$?"; tail -12 gpurun_out/pytest_dft.log
timeout 300 python tools/fft_bench.py > gpurun_out/fft_bench_dft.log 2>&1; echo "fft bench dft rc=$?"; cat gpurun_out/fft_bench_dft.log
ISDF_FFT_DFT=0 timeout 300 python tools/fft_bench.py > gpurun_out/fft_bench_stockham.log 2>&1; echo "fft bench stockham rc=$?"; cat gpurun_out/fft_bench_stockham.log
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity > gpurun_out/bench_c64_dft.json 2> gpurun_out/bench_c64_dft.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_c64_dft.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_c64_dft.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'])
for k,v in d['stages_ms_each'].items(): print(k, [round(x,1) for x in v])
PY
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dft_ -c 3 -f -o gpurun_out/prof_dft \
    python tools/fft_bench.py 71 > gpurun_out/ncu_dft.log 2>&1; echo "ncu dft rc=$?"; tail -2 gpurun_out/ncu_dft.log
