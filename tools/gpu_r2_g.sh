#!/bin/bash
# Round 2, call G (2 GPUs): suite with the new defaults (push pivot kernel, two-stream residue product, occupied-orbital J/K),
# C64 parity prefix, fft microbench on the sweep meshes, bench at N = 1 and N = 2 as the driver launches it.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
timeout 900 python tools/c64_parity.py --oracle-points 2048 --rows 0 --skip-planes > gpurun_out/c64_parity_newdefaults.json 2> gpurun_out/c64_parity_newdefaults.err; echo "parity rc=$?"; tail -1 gpurun_out/c64_parity_newdefaults.json
echo DFT; timeout 300 python tools/fft_bench.py 48 64 96 128 2>&1 | tee gpurun_out/fft_bench_pow2_dft.log
echo STOCKHAM; ISDF_FFT_DFT=0 timeout 300 python tools/fft_bench.py 48 64 96 128 2>&1 | tee gpurun_out/fft_bench_pow2_stockham.log
timeout 900 python bench.py --steps 3 --warmup 3 --no-cpu-baseline > gpurun_out/bench_n1.json 2> gpurun_out/bench_n1.err; echo "bench n1 rc=$?"; tail -2 gpurun_out/bench_n1.err
timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/bench_n2.json 2> gpurun_out/bench_n2.err; echo "bench n2 rc=$?"; tail -2 gpurun_out/bench_n2.err
ISDF_PROFILE_SELECT=1 timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29512 bench.py --gpus 2 --steps 2 --warmup 2 --no-e2e > gpurun_out/bench_n2_prof.json 2> gpurun_out/bench_n2_prof.err; echo "bench n2 prof rc=$?"
OMP_NUM_THREADS=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29513 bench.py --impl reference --gpus 2 --steps 1 --warmup 0 > gpurun_out/bench_ref_n2.json 2> gpurun_out/bench_ref_n2.err; echo "ref n2 rc=$?"; tail -c 600 gpurun_out/bench_ref_n2.json
python - <<'PY'
import json
for f in ('bench_n1','bench_n2','bench_n2_prof'):
    try:
        d=json.loads(open('gpurun_out/%s.json'%f).read().strip().splitlines()[-1])
        print(f, 'value', round(d['value'],4), 'e2e', d['e2e']['value'], 'jk', d['jk_s'], 'roof', round(d['roofline']['frac'],3))
        print('   ', {k: round(v,1) for k,v in d['stages_ms'].items()})
        if 'parity' in d: print('   parity', d['parity'])
    except Exception as e: print(f, 'no line', e)
PY
