#!/bin/bash
mkdir -p gpurun_out
timeout 300 python tools/syrk_i8_check.py 4096x65536 12480x131072 12480x362952 > gpurun_out/syrk_i8_check2.txt 2>&1; echo "check rc=$?"; tail -12 gpurun_out/syrk_i8_check2.txt
