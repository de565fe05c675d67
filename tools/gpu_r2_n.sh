#!/bin/bash
# Round 2, call N: GPU suite with the in-place congruence; outlier steps with the interpreter's GC off / on, 14 steps each.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
for gc in 0 1 0 1; do
  ISDF_BENCH_GC=$gc timeout 900 python bench.py --steps 14 --warmup 3 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_gc$gc.json 2> gpurun_out/bench_gc$gc.err; echo "gc $gc rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_gc$gc.json').read().strip().splitlines()[-1])
print('gc=$gc value', round(d['value'],4), 'wsolve', round(d['stages_ms']['wsolve_ms'],1))
print('  step ms ', [round(x) for x in d['each_step_ms']])
print('  select  ', [round(x) for x in d['stages_ms_each']['select_ms']])
print('  wgemm   ', [round(x) for x in d['stages_ms_each']['wgemm_ms']])
PY
done
