#!/bin/bash
# whole C64 step with the other CTA tiles of the TMA GEMM (sustained, power-capped regime: less L2 -> SM traffic per flop)
mkdir -p gpurun_out
for t in 0 8 7; do
  ISDF_GEMM_TILE=$t timeout 300 python bench.py --steps 3 --warmup 2 --no-e2e --no-cpu-baseline --no-parity --no-extra > gpurun_out/bench_tile$t.json 2> gpurun_out/bench_tile$t.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_tile$t.json').read().strip().splitlines()[-1])
print('tile $t value', round(d['value'],4), 'select', round(d['stages_ms']['select_ms'],1), 'wsolve', round(d['stages_ms']['wsolve_ms'],1), 'frac', round(d['roofline']['frac'],3), 'clk', d['clocks'].get('sm_mhz_mean_each_step'), 'W', d['clocks'].get('power_w_median_under_load'))
PY
done
