#!/bin/bash
# Round 2, call L: the contract line of the current code, 10 steps, with the clock distribution under load.
mkdir -p gpurun_out
timeout 1200 python bench.py --steps 10 --warmup 3 > gpurun_out/bench_contract.json 2> gpurun_out/bench_contract.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_contract.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_contract.json').read().strip().splitlines()[-1])
print('value', round(d['value'],4), 'e2e', d['e2e'], 'jk', d['jk_s'])
print([round(x) for x in d['each_step_ms']])
for k in ('select_ms','fft_ms','wgemm_ms','wsolve_ms'): print('   ', k, [round(x) for x in d['stages_ms_each'][k]])
print('clocks', d['clocks'])
print('roofline', {k: d['roofline'][k] for k in ('achieved','peak','frac','traffic')})
print('extra', d.get('extra'))
print('parity', d.get('parity'))
print('cpu', d['cpu_baseline']['value'], d['cpu_baseline']['stages_s'])
PY
