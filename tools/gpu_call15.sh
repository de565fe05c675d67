#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/syrk_i8_check5.txt
nvidia-smi --query-gpu=clocks.sm,power.draw,clocks_throttle_reasons.active --format=csv,noheader -lms 200 > gpurun_out/smi_syrk.txt &
SMI=$!
for cl in 4 1; do
  echo "== cluster $cl" >> gpurun_out/syrk_i8_check5.txt
  ISDF_SYRK_CLUSTER=$cl timeout 200 python tools/syrk_i8_check.py 12480x131072 12480x362952 >> gpurun_out/syrk_i8_check5.txt 2>&1; echo "cl=$cl rc=$?"
done
kill $SMI
cat gpurun_out/syrk_i8_check5.txt | grep -v "^$" | tail -30
sort gpurun_out/smi_syrk.txt | uniq -c | sort -rn | head -12
