#!/bin/bash
# Round 2, call K: step-time noise with the build work buffers kept on the object: 3 x 10 steps; GPU suite.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for run in a b c; do
  timeout 600 python bench.py --steps 10 --warmup 3 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_kept_$run.json 2> gpurun_out/bench_kept_$run.err; echo "run $run rc=$?"
done
python - <<'PY'
import json
for f in ('bench_kept_a','bench_kept_b','bench_kept_c'):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'],4), [round(x) for x in d['each_step_ms']])
        for k in ('select_ms','fft_ms','wgemm_ms','wsolve_ms'): print('   ', k, [round(x) for x in d['stages_ms_each'][k]])
    except Exception as e: print(f, 'no line', e)
PY
