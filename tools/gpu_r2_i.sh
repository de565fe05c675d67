#!/bin/bash
# Round 2, call I: suite; the bench line with the water box (config 3) as a one-GPU extra (aliased build buffers);
# C64 with the aliased layout forced, for its cost.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-parity > gpurun_out/bench_n1_extra.json 2> gpurun_out/bench_n1_extra.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_n1_extra.err
ISDF_ALIAS_ARENAS=1 timeout 600 python bench.py --steps 2 --warmup 2 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_n1_alias.json 2> gpurun_out/bench_n1_alias.err; echo "bench alias rc=$?"
python - <<'PY'
import json
for f in ('bench_n1_extra','bench_n1_alias'):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'],4), 'e2e', d['e2e']['value'])
        print('   ', {k: round(v,1) for k,v in d['stages_ms'].items()})
        for k,v in d.get('extra',{}).items(): print('   extra', k, v)
    except Exception as e: print(f, 'no line', e)
PY
nvidia-smi --query-gpu=memory.total,memory.used --format=csv
