#!/bin/bash
# Round 2, call R: FP64 residue reduction in the split kernel and two points per thread in the sweep: tests + C64 stages.
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_g1_stages_gpu.py tests/test_g2_syrk_i8_gpu.py tests/test_g3_isdf_gpu.py -m gpu -q -x > gpurun_out/pytest_part.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_part.log
timeout 600 python bench.py --steps 5 --warmup 2 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_r.json 2> gpurun_out/bench_r.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_r.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_r.json').read().strip().splitlines()[-1])
print('value', round(d['value'],4), [round(x) for x in d['each_step_ms']])
for k in ('select_ms','fft_ms','wgemm_ms','wsolve_ms'): print('   ', k, [round(x) for x in d['stages_ms_each'][k]])
PY
timeout 300 python tools/syrk_i8_check.py 12480x32768 2>&1 | tail -2
