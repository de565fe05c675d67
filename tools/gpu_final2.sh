#!/bin/bash
# End of round 2: what the driver runs -- GPU suite, smoke(), the default bench line, the reference arm briefly.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=5 > gpurun_out/pytest_gpu_final.log 2>&1; echo "pytest rc=$?"; tail -9 gpurun_out/pytest_gpu_final.log
timeout 300 python -c "import __graft_entry__ as g; g.smoke()" > gpurun_out/smoke_final.log 2>&1; echo "smoke rc=$?"; tail -4 gpurun_out/smoke_final.log
timeout 900 python bench.py > gpurun_out/bench_final_n1.json 2> gpurun_out/bench_final_n1.err; echo "bench rc=$?"; tail -2 gpurun_out/bench_final_n1.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_final_n1.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e'], 'launches', d['gpu_launches'], 'frac', d['roofline']['frac'])
print('each', d['each_step_ms'], 'clocks', d['clocks']['sm_mhz'], d['clocks'].get('sm_mhz_mean_each_step'))
print('parity', d['parity']); print('cpu', d['cpu_baseline']['value']); print('h2o32', d['extra']['h2o32']['value'])
PY
