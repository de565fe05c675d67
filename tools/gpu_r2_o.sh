#!/bin/bash
# Round 2, call O: which part of the selection carries an outlier step?  16 steps with the per-panel breakdown.
mkdir -p gpurun_out
ISDF_PROFILE_SELECT=1 timeout 900 python bench.py --steps 16 --warmup 3 --no-cpu-baseline --no-parity --no-extra --no-e2e > gpurun_out/bench_selprof.json 2> gpurun_out/bench_selprof.err; echo "rc=$?"
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_selprof.json').read().strip().splitlines()[-1])
print('step ms ', [round(x) for x in d['each_step_ms']])
for k in ('select_ms','sel_candidates_ms','sel_allgather_ms','sel_panel_ms','sel_flush_ms','wgemm_ms'):
    print('  %-18s' % k, [round(x) for x in d['stages_ms_each'][k]])
PY
