#!/bin/bash
# First GPU call of round 2.  State at the end of round 1 (profiles/r01_optin_forms.md): the residue form of the
# int8 metric product is the default (first runs passed, C64 wgemm 462 vs 737 ms); the cluster pivot kernels
# (ISDF_PANEL_CLUSTER=1/2) and the fused FFT (ISDF_FFT_FUSED=1) passed their first tests and stay opt-in.
# This call: (1) the whole GPU suite under the new default, (2) the cluster pivot forms again plus
# compute-sanitizer over the late kernels, (3) C64 bench lines per pivot form with the selection breakdown,
# (4) fused FFT line, (5) the digit-plane product for comparison (ISDF_SYRK_CRT=0), (6) nr_rks('pbe') timing.
mkdir -p gpurun_out
timeout 1200 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -3 gpurun_out/pytest_gpu.log
ISDF_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_zz_panel_forms_gpu.py -m gpu -q > gpurun_out/pytest_forms.log 2>&1; echo "forms rc=$?"; tail -15 gpurun_out/pytest_forms.log
ISDF_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_zz_fft_fused_gpu.py -m gpu -q > gpurun_out/pytest_fftfused.log 2>&1; echo "fft fused rc=$?"; tail -15 gpurun_out/pytest_fftfused.log
ISDF_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_zz_xc_gpu.py tests/test_zz_scf_gpu.py -m gpu -q > gpurun_out/pytest_xc.log 2>&1; echo "xc rc=$?"; tail -15 gpurun_out/pytest_xc.log
# memory checker on the small experimental tests (slow: only these files)
ISDF_TEST_EXPERIMENTAL=1 timeout 900 compute-sanitizer --tool memcheck --error-exitcode 9 python -m pytest \
    tests/test_zz_panel_forms_gpu.py tests/test_g2_syrk_i8_gpu.py tests/test_zz_xc_gpu.py -m gpu -q -x -k "bit_exact or 200-300 or 128-128 or he_c or lda" \
    > gpurun_out/sanitizer.log 2>&1; echo "sanitizer rc=$?"; grep -E "ERROR SUMMARY|passed|failed" gpurun_out/sanitizer.log | tail -5
for mode in 0 1 2; do
  ISDF_PANEL_CLUSTER=$mode ISDF_PROFILE_SELECT=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e \
      > gpurun_out/bench_panel$mode.json 2> gpurun_out/bench_panel$mode.err; echo "bench mode $mode rc=$?"
  python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_panel$mode.json').read().strip().splitlines()[-1])
    print('mode $mode step %.3f s' % d['value'], {k: round(v, 1) for k, v in d['stages_ms'].items()})
except Exception as e:
    print('mode $mode: no line', e)
PY
done
ISDF_FFT_FUSED=1 ISDF_PROFILE_SELECT=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e \
    > gpurun_out/bench_fftfused.json 2> gpurun_out/bench_fftfused.err; echo "bench fft fused rc=$?"
python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_fftfused.json').read().strip().splitlines()[-1])
    print('fft fused step %.3f s' % d['value'], {k: round(v, 1) for k, v in d['stages_ms'].items()})
except Exception as e:
    print('fft fused: no line', e)
PY
ISDF_SYRK_CRT=0 ISDF_SYRK_VERBOSE=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e \
    > gpurun_out/bench_crt.json 2> gpurun_out/bench_crt.err; echo "bench crt rc=$?"; grep -m2 isdf_syrk gpurun_out/bench_crt.err
python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_crt.json').read().strip().splitlines()[-1])
    print('crt step %.3f s' % d['value'], {k: round(v, 1) for k, v in d['stages_ms'].items()}, d.get('wgemm_i8'))
except Exception as e:
    print('crt: no line', e)
PY
# J/K + exchange-correlation per SCF iteration for the RKS-PBE configuration (BASELINE configs[1])
timeout 600 python bench.py --steps 1 --warmup 1 --no-cpu-baseline --no-e2e --xc pbe \
    > gpurun_out/bench_xc.json 2> gpurun_out/bench_xc.err; echo "bench xc rc=$?"
python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_xc.json').read().strip().splitlines()[-1])
    print('xc', d.get('xc'), 'jk_ms', d.get('jk_ms'))
except Exception as e:
    print('xc: no line', e)
PY
# two-stream form of the residue product (ISDF_SYRK_OVERLAP=1): bit-identity test, then a C64 line
ISDF_TEST_EXPERIMENTAL=1 timeout 300 python -m pytest tests/test_g2_syrk_i8_gpu.py -m gpu -q -k overlapped > gpurun_out/pytest_overlap.log 2>&1; echo "overlap rc=$?"; tail -5 gpurun_out/pytest_overlap.log
ISDF_SYRK_OVERLAP=1 timeout 600 python bench.py --steps 2 --warmup 1 --no-cpu-baseline --no-e2e \
    > gpurun_out/bench_overlap.json 2> gpurun_out/bench_overlap.err; echo "bench overlap rc=$?"
python - <<PY
import json
try:
    d = json.loads(open('gpurun_out/bench_overlap.json').read().strip().splitlines()[-1])
    print('overlap step %.3f s' % d['value'], {k: round(v, 1) for k, v in d['stages_ms'].items()})
except Exception as e:
    print('overlap: no line', e)
PY
