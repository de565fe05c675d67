#!/bin/bash
mkdir -p gpurun_out
export ISDF_PROFILE_SELECT=1
n=2; w=c64
timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --workload $w --steps 2 --warmup 1 > gpurun_out/scale_${w}_n$n.json 2> gpurun_out/scale_${w}_n$n.err
echo "$w N=$n rc=$?"; grep -v "^$" gpurun_out/scale_${w}_n$n.err | grep -iv "warn\|\*\*\*\|OMP_NUM" | tail -3
python -c "
import json
d=json.loads(open('gpurun_out/scale_${w}_n$n.json').read().strip().splitlines()[-1]); print('$w N=$n', d['value'], d['build_s'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d['roofline']['achieved'])"
