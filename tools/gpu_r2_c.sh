#!/bin/bash
# Round 2, call C: the bench line under the contract (new e2e / parity / cpu legs), the reference arm, then the
# full C64 parity check (all 12480 pivots against the blocked oracle: ~5 min of host FMA chains).
mkdir -p gpurun_out
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/bench_c64.json 2> gpurun_out/bench_c64.err; echo "bench rc=$?"; tail -c 3000 gpurun_out/bench_c64.json; tail -5 gpurun_out/bench_c64.err
OMP_NUM_THREADS=1 timeout 900 python bench.py --impl reference --steps 2 --warmup 1 > gpurun_out/bench_ref.json 2> gpurun_out/bench_ref.err; echo "ref rc=$?"; tail -c 2500 gpurun_out/bench_ref.json; tail -5 gpurun_out/bench_ref.err
timeout 1500 python tools/c64_parity.py --skip-planes --rows 0 > gpurun_out/c64_parity_full.json 2> gpurun_out/c64_parity_full.err; echo "c64 parity rc=$?"; tail -1 gpurun_out/c64_parity_full.json; tail -5 gpurun_out/c64_parity_full.err
