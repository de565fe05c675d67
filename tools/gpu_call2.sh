#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
tail -15 gpurun_out/pytest_gpu.log
for t in 0 1 2; do
ISDF_GEMM_TILE=$t timeout 900 python bench.py --workload c64 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c64_tile$t.json 2> gpurun_out/bench_c64_tile$t.err; echo "c64 tile$t rc=$?"
done
timeout 600 python bench.py --workload c8 --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_c8.json 2> gpurun_out/bench_c8.err; echo "c8 rc=$?"
