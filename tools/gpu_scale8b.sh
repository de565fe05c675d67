#!/bin/bash
# 8 GPUs, second pass: the driver's launch line again (Si216 extra after the workspace fix), reference arm under torchrun.
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench n8 rc=$?"; tail -3 gpurun_out/bench_n8.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_n8.json').read().strip().splitlines()[-1])
print('value', round(d['value'],4), 'e2e', d['e2e']['value'], 'jk', d['jk_s'])
for k,v in d.get('extra',{}).items(): print('   extra', k, v)
PY
