#!/bin/bash
# Round 2, call J: where do the sporadic ~1 s stalls of a step come from?  Ten steps each with the nvidia-smi
# sampler off / at 1000 ms / at 200 ms; then the aliased layout again (row alignment fixed).
mkdir -p gpurun_out
for ms in 0 1000 200; do
  ISDF_BENCH_SMI_MS=$ms timeout 600 python bench.py --steps 10 --warmup 2 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_smi$ms.json 2> gpurun_out/bench_smi$ms.err; echo "smi $ms rc=$?"
done
ISDF_ALIAS_ARENAS=1 timeout 600 python bench.py --steps 4 --warmup 2 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_n1_alias.json 2> gpurun_out/bench_n1_alias.err; echo "bench alias rc=$?"
python - <<'PY'
import json
for f in ('bench_smi0','bench_smi1000','bench_smi200','bench_n1_alias'):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'],4), [round(x) for x in d['each_step_ms']])
        for k in ('select_ms','fft_ms','wgemm_ms'): print('   ', k, [round(x) for x in d['stages_ms_each'][k]])
    except Exception as e: print(f, 'no line', e)
PY
