#!/bin/bash
# Round 2: the residue (default) form of the int8 metric product under ncu -- one chunk of the C64 shape
# (n = 12480, K = 32768): launch times of the three kernels, then one --set full capture of each.
# Run only after `python tools/syrk_i8_check.py 12480x32768` has exited 0 without ncu.
mkdir -p gpurun_out
timeout 300 python tools/syrk_i8_check.py 12480x32768 > gpurun_out/syrk_crt_plain.log 2>&1; echo "plain rc=$?"; tail -3 gpurun_out/syrk_crt_plain.log
timeout 600 ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__pipe_tensor_cycles_active.avg.pct_of_peak_sustained_elapsed \
    --clock-control none -k regex:"syrk_i8_kernel2|i8_split_crt|crt_reconstruct|i8_rowstats" -c 12 --csv --log-file gpurun_out/ncu_syrk_crt_launches.csv \
    python tools/syrk_i8_check.py 12480x32768 > gpurun_out/ncu_syrk_crt_launches.log 2>&1; echo "ncu launches rc=$?"
for k in syrk_i8_kernel2 i8_split_crt_kernel crt_reconstruct_kernel; do
  timeout 900 ncu --set full --clock-control none --import-source on -k regex:$k -c 1 -o gpurun_out/ncu_$k \
      python tools/syrk_i8_check.py 12480x32768 > gpurun_out/ncu_$k.log 2>&1; echo "ncu $k rc=$?"
done
