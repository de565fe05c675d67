#!/bin/bash
# 2 GPUs: the driver's launch line (C64 headline + the water box as an extra), then the reference arm as launched at N = 2.
mkdir -p gpurun_out
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -2 gpurun_out/bench_n2.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n2.json').read().strip().splitlines()[-1])
print('value', round(d['value'],4), 'e2e', d['e2e']['value'], 'jk', d['jk_s'])
print('   ', {k: round(v,1) for k,v in d['stages_ms'].items()})
for k,v in d.get('extra',{}).items(): print('   extra', k, v)
PY
