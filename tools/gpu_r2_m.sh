#!/bin/bash
# Round 2, call M: do the slow steps coincide with low SM clocks?  14 steps, per-step mean clock in the bench line.
mkdir -p gpurun_out
timeout 900 python bench.py --steps 14 --warmup 3 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_clock_corr.json 2> gpurun_out/bench_clock_corr.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_clock_corr.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_clock_corr.json').read().strip().splitlines()[-1])
print('value', round(d['value'],4))
print('step ms ', [round(x) for x in d['each_step_ms']])
print('select  ', [round(x) for x in d['stages_ms_each']['select_ms']])
print('wgemm   ', [round(x) for x in d['stages_ms_each']['wgemm_ms']])
print('mean MHz', [None if x is None else round(x) for x in d['clocks']['sm_mhz_mean_each_step']])
print({k:v for k,v in d['clocks'].items() if k != 'sm_mhz_mean_each_step'})
PY
