#!/bin/bash
# 8 GPUs with the TMA-fed GEMM: the driver's launch line without the one-off extras (budget), then N = 4 on the same box
mkdir -p gpurun_out
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29531 bench.py --gpus 8 --steps 3 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/bench_n8_tma.json 2> gpurun_out/bench_n8_tma.err; echo "bench n8 rc=$?"; tail -2 gpurun_out/bench_n8_tma.err
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 4 --master-addr 127.0.0.1 --master-port 29532 bench.py --gpus 4 --steps 3 --warmup 3 --no-extra --no-cpu-baseline > gpurun_out/bench_n4_tma.json 2> gpurun_out/bench_n4_tma.err; echo "bench n4 rc=$?"; tail -2 gpurun_out/bench_n4_tma.err
python - <<'PY'
import json
for n in (8, 4):
    d=json.loads(open('gpurun_out/bench_n%d_tma.json' % n).read().strip().splitlines()[-1])
    print(n, 'value', round(d['value'],4), 'e2e', d['e2e'], 'jk', d['jk_s'], 'frac', round(d['roofline']['frac'],3), d['stages_ms'])
PY
