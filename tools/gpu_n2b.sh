#!/bin/bash
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 3 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/bench_n2_tma.json 2> gpurun_out/bench_n2_tma.err; echo "bench n2 rc=$?"; tail -2 gpurun_out/bench_n2_tma.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n2_tma.json').read().strip().splitlines()[-1])
print(2, 'value', round(d['value'],4), 'e2e', d['e2e']['value'], 'build', d['build_s'], 'jk', d['jk_s'], 'frac', round(d['roofline']['frac'],3), d['stages_ms'])
PY
