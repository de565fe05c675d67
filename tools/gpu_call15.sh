#!/bin/bash
mkdir -p gpurun_out
: > gpurun_out/syrk_i8_check6.txt
for kv in 2 1; do
  echo "== kernel variant $kv" >> gpurun_out/syrk_i8_check6.txt
  ISDF_SYRK_KERNEL=$kv timeout 200 python tools/syrk_i8_check.py 128x128 200x300 700x5000 1000x70001 4096x65536 12480x131072 >> gpurun_out/syrk_i8_check6.txt 2>&1; echo "kv=$kv rc=$?"
done
cat gpurun_out/syrk_i8_check6.txt | grep -v "^$" | tail -30
