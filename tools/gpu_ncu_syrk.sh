#!/bin/bash
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:syrk_i8_kernel -c 1 -f -o gpurun_out/prof_syrk_i8_final python tools/syrk_i8_check.py 12480x32768 > gpurun_out/ncu_syrk.log 2>&1; echo "ncu rc=$?"; tail -3 gpurun_out/ncu_syrk.log
