#!/bin/bash
# blocked panel sweep (earlier row blocks through the FP64 GEMM): GPU suite, bit-equality of the new 32x128 tile, step time per block size
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_sweep.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_sweep.log
timeout 200 python tools/gemm_tma_check.py > gpurun_out/gemm_tma_check2.log 2>&1; echo "tma check rc=$?"; tail -2 gpurun_out/gemm_tma_check2.log
for b in 128 64 32 16; do
  ISDF_SWEEP_BLOCK=$b timeout 300 python bench.py --steps 2 --warmup 2 --no-e2e --no-cpu-baseline --no-extra > gpurun_out/bench_sweep$b.json 2> gpurun_out/bench_sweep$b.err
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_sweep$b.json').read().strip().splitlines()[-1])
print('block $b value', round(d['value'],4), 'select', round(d['stages_ms']['select_ms'],1), 'frac', round(d['roofline']['frac'],3), 'parity', d['parity']['ip_equal'], d['parity']['resid_equal'], d['parity']['max_dJ'])
PY
done
