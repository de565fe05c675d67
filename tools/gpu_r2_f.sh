#!/bin/bash
# Round 2, call F: full suite with the dense-DFT passes, experimental forms, C64 bench line, sanitizer record, sweep.
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -m gpu -q -x > gpurun_out/pytest_gpu.log 2>&1; echo "pytest rc=$?"; tail -4 gpurun_out/pytest_gpu.log
ISDF_TEST_EXPERIMENTAL=1 timeout 600 python -m pytest tests/test_zz_panel_forms_gpu.py tests/test_g2_syrk_i8_gpu.py tests/test_zz_fft_fused_gpu.py -m gpu -q > gpurun_out/pytest_forms.log 2>&1; echo "forms rc=$?"; tail -4 gpurun_out/pytest_forms.log
timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity > gpurun_out/bench_c64_dft.json 2> gpurun_out/bench_c64_dft.err; echo "bench rc=$?"; tail -3 gpurun_out/bench_c64_dft.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_c64_dft.json').read().strip().splitlines()[-1])
print('value', d['value'], 'e2e', d['e2e']['value'], 'jk', d['jk_s'])
for k,v in d['stages_ms_each'].items(): print(k, [round(x,1) for x in v])
PY
ISDF_PANEL_CLUSTER=2 ISDF_SYRK_OVERLAP=1 timeout 600 python bench.py --steps 3 --warmup 2 --no-cpu-baseline --no-parity --no-e2e > gpurun_out/bench_c64_forms.json 2> gpurun_out/bench_c64_forms.err; echo "bench forms rc=$?"; tail -3 gpurun_out/bench_c64_forms.err
python - <<'PY'
import json
d=json.loads(open('gpurun_out/bench_c64_forms.json').read().strip().splitlines()[-1])
print('push+overlap value', d['value'])
for k,v in d['stages_ms_each'].items(): print(k, [round(x,1) for x in v])
PY
SAN="tests/test_g1_stages_gpu.py tests/test_g2_syrk_i8_gpu.py tests/test_zz_panel_forms_gpu.py"
SEL="select_ip_bit_exact or forward_passes_match_numpy_fft or three_threads or 257-1000 or cluster_forms or build_W"
for tool in memcheck racecheck synccheck; do
  ISDF_TEST_EXPERIMENTAL=1 timeout 1200 compute-sanitizer --tool $tool --error-exitcode 9 python -m pytest $SAN -m gpu -q -x -k "$SEL" \
      > gpurun_out/sanitizer_$tool.log 2>&1; echo "sanitizer $tool rc=$?"; grep -E "ERROR SUMMARY|RACECHECK SUMMARY|passed|failed" gpurun_out/sanitizer_$tool.log | tail -4
done
timeout 900 python tools/sweep.py --out gpurun_out/sweep.json > gpurun_out/sweep.log 2>&1; echo "sweep rc=$?"; tail -22 gpurun_out/sweep.log
