#!/bin/bash
# k_convolve: producer / epilogue waits sleep between polls (A/B against the tight loops), small parity run
mkdir -p gpurun_out; : > gpurun_out/s21_ab.txt
NO_ORACLE=1 timeout 300 python tools/conv_ab.py sleep c1 c1b c4 >> gpurun_out/s21_ab.txt 2>&1; echo "sleep rc $?" >> gpurun_out/s21_ab.txt
ZB_LIB=build/libzb200_spin.so NO_ORACLE=1 timeout 300 python tools/conv_ab.py spin c1 c1b c4 >> gpurun_out/s21_ab.txt 2>&1; echo "spin rc $?" >> gpurun_out/s21_ab.txt
cut -c1-160 gpurun_out/s21_ab.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "small_cases_host or full_size_quadrant" > gpurun_out/s21_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s21_pytest.txt; tail -3 gpurun_out/s21_pytest.txt | cut -c1-200
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/s21_bench.json 2> gpurun_out/s21_bench.err; echo "bench rc $?"
python - <<'PY'
import json
r=json.loads([l for l in open('gpurun_out/s21_bench.json') if l.startswith('{')][-1])
print(r['value'], r['e2e']['value'], r['stage_ms_serial'])
PY
