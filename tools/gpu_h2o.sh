#!/bin/bash
mkdir -p gpurun_out
export ISDF_PROFILE_SELECT=1
n=4; w=h2o32
timeout 500 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --workload $w --steps 1 --warmup 1 > gpurun_out/scale_${w}_n$n.json 2> gpurun_out/scale_${w}_n$n.err
echo "$w N=$n rc=$?"; grep -v "^$" gpurun_out/scale_${w}_n$n.err | grep -iv "warn\|\*\*\*\|OMP_NUM" | tail -3
python -c "
import json
d=json.loads(open('gpurun_out/scale_${w}_n$n.json').read().strip().splitlines()[-1]); print('$w N=$n', d['value'], d['build_s'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d['roofline']['achieved'], d['config'], d.get('wgemm_i8'))"
