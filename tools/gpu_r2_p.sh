#!/bin/bash
# Round 2, call P: host-side facts during the timed loop (CFS throttling, context switches, CPU time) for 16 steps,
# once as is and once with the host math libraries held to one thread.
mkdir -p gpurun_out
cat /sys/fs/cgroup/cpu.max /sys/fs/cgroup/cpu.stat 2>&1 | head -12; nproc
for mode in default onethread; do
  if [ $mode = onethread ]; then export OMP_NUM_THREADS=1 OPENBLAS_NUM_THREADS=1 MKL_NUM_THREADS=1; fi
  timeout 900 python bench.py --steps 16 --warmup 3 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_host_$mode.json 2> gpurun_out/bench_host_$mode.err; echo "$mode rc=$?"
  python - <<PY
import json
d=json.loads(open('gpurun_out/bench_host_$mode.json').read().strip().splitlines()[-1])
print('$mode value', round(d['value'],4))
print('  step ms ', [round(x) for x in d['each_step_ms']])
print('  select  ', [round(x) for x in d['stages_ms_each']['select_ms']])
print('  host', d['host'])
PY
done
