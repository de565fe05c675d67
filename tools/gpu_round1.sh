#!/bin/bash
# first GPU pass: parity tests, FP64 probe, bench (c8 + c64), ncu launch list
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,memory.total,clocks.max.sm --format=csv > gpurun_out/gpu.txt 2>&1
./tools/fp64_probe > gpurun_out/fp64_probe.txt 2>&1
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?" >> gpurun_out/pytest_gpu.log
timeout 600 python bench.py --workload c8 --steps 3 --warmup 3 > gpurun_out/bench_c8.json 2> gpurun_out/bench_c8.err; echo "c8 rc=$?"
timeout 1200 python bench.py --workload c64 --steps 2 --warmup 1 > gpurun_out/bench_c64.json 2> gpurun_out/bench_c64.err; echo "c64 rc=$?"
timeout 600 python bench.py --workload c8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/plain.log 2>&1 &&
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches_c8.csv python bench.py --workload c8 --steps 1 --warmup 1 --no-cpu-baseline > gpurun_out/ncu_c8.log 2>&1
echo "ncu rc=$?"
tail -3 gpurun_out/pytest_gpu.log
cat gpurun_out/fp64_probe.txt
