#!/bin/bash
# usage: gpu_multi.sh "<N list>" workload steps warmup
mkdir -p gpurun_out
nvidia-smi --query-gpu=index,name --format=csv,noheader | head -8
export ISDF_PROFILE_SELECT=1
for n in $1; do
  if [ "$n" = "1" ]; then
    timeout 400 python bench.py --gpus 1 --workload $2 --steps $3 --warmup $4 --no-cpu-baseline > gpurun_out/scale_$2_n$n.json 2> gpurun_out/scale_$2_n$n.err
  else
    timeout 400 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 29533 bench.py --gpus $n --workload $2 --steps $3 --warmup $4 > gpurun_out/scale_$2_n$n.json 2> gpurun_out/scale_$2_n$n.err
  fi
  echo "N=$n rc=$?"; grep -v "^$" gpurun_out/scale_$2_n$n.err | grep -iv "warn\|\*\*\*\|OMP_NUM" | tail -4
  python -c "
import json
d=json.loads(open('gpurun_out/scale_$2_n$n.json').read().strip().splitlines()[-1]); print('N=$n', d['value'], d['build_s'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d['roofline']['achieved'])"
done
