#!/bin/bash
# Round 2, call H: full GPU suite (device non-local pp, occupied-orbital J/K), the launch list of one C64 step,
# ncu --set full of the flush GEMM (traffic figure of the roofline object) and of the residue-product kernels.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x --durations=6 > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -14 gpurun_out/pytest_gpu.log
timeout 300 python - <<'PY' > gpurun_out/pp_nl_timing.log 2>&1
import time, numpy as np, torch
from pyscf_isdf_b200 import workloads
from pyscf_isdf_b200.isdf import ISDF
cell = workloads.make_cell('c64')
mydf = ISDF(cell, c=15)
for dev in (True, True, False):
    t0 = time.perf_counter(); v = mydf.get_pp_nl(on_device=dev); torch.cuda.synchronize(); print('get_pp_nl on_device=%s: %.3f s  fp %.12f' % (dev, time.perf_counter() - t0, float(abs(v).sum())), flush=True)
t0 = time.perf_counter(); v = mydf.get_pp_loc(); print('get_pp_loc: %.3f s' % (time.perf_counter() - t0))
PY
cat gpurun_out/pp_nl_timing.log
timeout 900 ncu --metrics gpu__time_duration.sum --clock-control none -c 20000 --csv --log-file gpurun_out/launches_c64.csv \
    python bench.py --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-parity > gpurun_out/ncu_launches.log 2>&1; echo "launch list rc=$?"
python tools/summarize_launches.py gpurun_out/launches_c64.csv gpurun_out/launches_c64.md | head -30
timeout 600 ncu --set full --clock-control none --import-source on -k regex:dgemm_kernel -c 2 -f -o gpurun_out/prof_dgemm_r02 python tools/gemm_bench.py 0 > gpurun_out/ncu_dgemm.log 2>&1; echo "ncu gemm rc=$?"; tail -3 gpurun_out/ncu_dgemm.log
ISDF_SYRK_OVERLAP=0 bash tools/gpu_ncu_syrk_crt.sh
