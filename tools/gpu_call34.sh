#!/bin/bash
mkdir -p gpurun_out
export ISDF_SYRK_VERBOSE=1
timeout 600 python bench.py --workload c64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c64_k2.json 2> gpurun_out/bench_c64_k2.err; echo "c64 rc=$?"; tail -2 gpurun_out/bench_c64_k2.err
python -c "
import json,sys; d=json.loads(open('gpurun_out/bench_c64_k2.json').read().strip().splitlines()[-1]); print(d['value'], d['jk_s'], d['e2e']['value'], d['stages_ms'], d.get('wgemm_i8'), d['roofline']['achieved'], d['clocks'])"
unset ISDF_SYRK_VERBOSE
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,l1tex__m_xbar2l1tex_read_bytes.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.max.per_second --clock-control none -k regex:syrk_i8_kernel -c 1 python tools/syrk_i8_check.py 12480x32768 > gpurun_out/ncu_syrk_k2.log 2>&1; echo "ncu rc=$?"; grep -A12 "syrk_i8_kernel" gpurun_out/ncu_syrk_k2.log | tail -14
