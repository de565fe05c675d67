#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -5 gpurun_out/pytest_gpu.log
export ISDF_PROFILE_SELECT=1
export ISDF_SYRK_VERBOSE=1
for w in i8 i8:7; do
timeout 600 python bench.py --workload c64 --steps 2 --warmup 1 --no-cpu-baseline --wgemm $w > gpurun_out/bench_c64_$w.json 2> gpurun_out/bench_c64_$w.err; echo "c64 $w rc=$?"; tail -2 gpurun_out/bench_c64_$w.err
python -c "
import json,sys; d=json.loads(open('gpurun_out/bench_c64_$w.json').read().strip().splitlines()[-1]); print(d['value'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d.get('wgemm_i8'), d['clocks'])"
done
