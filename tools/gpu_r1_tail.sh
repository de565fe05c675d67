#!/bin/bash
# Last GPU seconds of round 1: the opt-in kernel forms written after the main GPU sessions -- tests, then one
# exploratory C64 bench line each (1 warm-up + 1 step: for comparing stage times, not a reportable number).
# Kept as the record of what ran (profiles/r01_optin_forms.md); tests/test_zz_syrk_crt_gpu.py was folded into
# tests/test_g2_syrk_i8_gpu.py afterwards, when the residue form became the default.
mkdir -p gpurun_out
B="python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e"
ISDF_TEST_EXPERIMENTAL=1 timeout 50 python -m pytest tests/test_zz_fft_fused_gpu.py tests/test_zz_syrk_crt_gpu.py -m gpu -q -rf -p no:cacheprovider > gpurun_out/r1_tail_tests1.log 2>&1; echo "tests fft-fused+crt rc=$?"; tail -4 gpurun_out/r1_tail_tests1.log
timeout 45 $B --xc pbe > gpurun_out/r1_tail_bench_xc.json 2> gpurun_out/r1_tail_bench_xc.err; echo "bench default+xc rc=$?"
ISDF_FFT_FUSED=1 timeout 40 $B > gpurun_out/r1_tail_bench_fftfused.json 2> gpurun_out/r1_tail_bench_fftfused.err; echo "bench fft fused rc=$?"
ISDF_SYRK_CRT=16 timeout 40 $B > gpurun_out/r1_tail_bench_crt.json 2> gpurun_out/r1_tail_bench_crt.err; echo "bench crt rc=$?"
ISDF_TEST_EXPERIMENTAL=1 timeout 40 python -m pytest tests/test_zz_panel_forms_gpu.py -m gpu -q -rf -p no:cacheprovider > gpurun_out/r1_tail_tests2.log 2>&1; echo "tests panel forms rc=$?"; tail -4 gpurun_out/r1_tail_tests2.log
ISDF_PANEL_CLUSTER=2 timeout 40 $B > gpurun_out/r1_tail_bench_panel2.json 2> gpurun_out/r1_tail_bench_panel2.err; echo "bench panel push rc=$?"
python - <<'PY'
import json, glob
for f in sorted(glob.glob('gpurun_out/r1_tail_bench_*.json')):
    try:
        d = json.loads(open(f).read().strip().splitlines()[-1])
        print(f.split('bench_')[1], 'step %.3f s' % d['value'], {k: round(v, 1) for k, v in d.get('stages_ms', {}).items()}, d.get('xc'))
    except Exception as e:
        print(f, 'no line', e)
PY
