#!/bin/bash
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name,memory.used --format=csv,noheader | head -8
export ISDF_PROFILE_SELECT=1
run() {  # n workload steps warmup extra
  local n=$1 w=$2
  timeout 420 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --workload $w --steps $3 --warmup $4 $5 > gpurun_out/scale_${w}_n$n.json 2> gpurun_out/scale_${w}_n$n.err
  echo "$w N=$n rc=$?"; grep -v "^$" gpurun_out/scale_${w}_n$n.err | grep -iv "warn\|\*\*\*\|OMP_NUM" | tail -3
  python -c "
import json
d=json.loads(open('gpurun_out/scale_${w}_n$n.json').read().strip().splitlines()[-1]); print('$w N=$n', d['value'], d['build_s'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d['roofline']['achieved'])"
}
run 8 c64 2 1 ""
run 4 c64 2 1 ""
# run 8 si216 1 1 "--no-e2e"
