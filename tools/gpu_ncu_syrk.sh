#!/bin/bash
mkdir -p gpurun_out
ISDF_SYRK_DEBUG_SAMETILE=1 timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,lts__t_sector_hit_rate.pct,l1tex__m_xbar2l1tex_read_bytes.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed,sm__cycles_elapsed.max.per_second --clock-control none -k regex:syrk_i8_kernel -c 1 python tools/syrk_i8_check.py 12480x32768 > gpurun_out/ncu_syrk_same.log 2>&1; echo "ncu rc=$?"; grep -A12 "syrk_i8_kernel" gpurun_out/ncu_syrk_same.log | tail -14
