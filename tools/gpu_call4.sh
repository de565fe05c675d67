#!/bin/bash
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
timeout 900 python bench.py --workload c64 --steps 2 --warmup 1 --no-cpu-baseline > gpurun_out/bench_c64.json 2> gpurun_out/bench_c64.err; echo "c64 rc=$?"
timeout 600 python bench.py --workload c64 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 1200 ncu --metrics gpu__time_duration.sum --clock-control none -c 30000 --csv --log-file gpurun_out/launches_c64.csv python bench.py --workload c64 --steps 1 --warmup 0 --no-cpu-baseline > gpurun_out/ncu_c64.log 2>&1
echo "ncu rc=$?"
