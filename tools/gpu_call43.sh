#!/bin/bash
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
export ISDF_PROFILE_SELECT=1
timeout 600 python bench.py --workload c64 --steps 3 --warmup 2 --no-cpu-baseline > gpurun_out/bench_c64_m.json 2> gpurun_out/bench_c64_m.err; echo "c64 rc=$?"; tail -2 gpurun_out/bench_c64_m.err
python -c "
import json,sys; d=json.loads(open('gpurun_out/bench_c64_m.json').read().strip().splitlines()[-1]); print(d['value'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d['roofline']['achieved'], d['roofline']['frac'], d['clocks'])"
