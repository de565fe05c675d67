#!/bin/bash
# protocol variants of the TMA GEMM (build_variants/lib_<variant>_<stages>.so) on the shapes that fail most often
for v in "$@"; do
  for mk in "64 192" "64 6240" "52 194"; do set -- $mk
    echo "== variant $v M $1 K $2: $(M=$1 K=$2 LIB=build_variants/lib_$v.so FORMS=10 REPS=${REPS:-200} timeout 160 python tools/gemm_tma_repro.py 2>&1 | tail -1)"
  done
  echo "   perf: $(GEMM_CASE=flush LIB=build_variants/lib_$v.so timeout 100 python tools/gemm_bench.py 0 2>&1 | tail -1 | cut -c60-140)"
done
