#!/bin/bash
# Protocol variants of the TMA GEMM on the shapes that fail most often (record: profiles/r02_gemm_tma.md).
# The variants were builds of csrc/gemm_tma.cu with -DTMA_VARIANT=<bits> -DTMA_NSTAGE=<n> at commit 56c8826
# (bit 0: arrive after the DMMAs, 1: fence.proxy.async in every lane before the arrive, 2: __syncwarp after the wait,
# 3: the fence in lane 0 only, 4: the fence after the wait), linked into build_variants/lib_<bits>_<n>.so; HEAD keeps
# only the form that passed (bit 1).  tools/gemm_tma_repro.py LIB=<path> runs any library against the LDGSTS kernel.
for v in "$@"; do
  for mk in "64 192" "64 6240" "52 194"; do set -- $mk
    echo "== variant $v M $1 K $2: $(M=$1 K=$2 LIB=build_variants/lib_$v.so FORMS=10 REPS=${REPS:-200} timeout 160 python tools/gemm_tma_repro.py 2>&1 | tail -1)"
  done
  echo "   perf: $(GEMM_CASE=flush LIB=build_variants/lib_$v.so timeout 100 python tools/gemm_bench.py 0 2>&1 | tail -1 | cut -c60-140)"
done
