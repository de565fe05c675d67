#!/bin/bash
# 8 GPUs: the driver's launch line (C64 headline + the one-off h2o32 / si216 runs under `extra`), then the
# selection / FFT breakdown of C64 (no extras).
mkdir -p gpurun_out
timeout 1500 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29521 bench.py --gpus 8 --steps 3 --warmup 3 > gpurun_out/bench_n8.json 2> gpurun_out/bench_n8.err; echo "bench n8 rc=$?"; tail -3 gpurun_out/bench_n8.err
ISDF_PROFILE_SELECT=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29522 bench.py --gpus 8 --steps 2 --warmup 2 --no-e2e --no-extra > gpurun_out/bench_n8_prof.json 2> gpurun_out/bench_n8_prof.err; echo "bench n8 prof rc=$?"
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29523 bench.py --gpus 4 --steps 3 --warmup 3 --no-extra > gpurun_out/bench_n4.json 2> gpurun_out/bench_n4.err; echo "bench n4 rc=$?"
python - <<'PY'
import json
for f in ('bench_n8','bench_n8_prof','bench_n4'):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'],4), 'e2e', d['e2e']['value'], 'jk', d['jk_s'], 'roof', round(d['roofline']['frac'],3))
        print('   ', {k: round(v,1) for k,v in d['stages_ms'].items()})
        for k,v in d.get('extra',{}).items(): print('   extra', k, v)
    except Exception as e: print(f, 'no line', e)
PY
