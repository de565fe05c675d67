#!/bin/bash
# Round 2, call Q: outlier steps against the host thread pools: 24 steps each, default / OMP=1 / OPENBLAS=1 / both.
mkdir -p gpurun_out
run() {
  timeout 900 python bench.py --steps 24 --warmup 3 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_pool_$1.json 2> gpurun_out/bench_pool_$1.err; echo "$1 rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_pool_$1.json').read().strip().splitlines()[-1])
ms=d['each_step_ms']; print('$1 value', round(d['value'],4), 'outliers(>3150)', sum(1 for x in ms if x>3150), [round(x) for x in ms])
PY
}
run default
OMP_NUM_THREADS=1 run omp1
OPENBLAS_NUM_THREADS=1 run openblas1
OMP_NUM_THREADS=1 OPENBLAS_NUM_THREADS=1 MKL_NUM_THREADS=1 run both1
run default2
