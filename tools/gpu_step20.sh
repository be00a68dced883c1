#!/bin/bash
# k_fill_conv: float rounding of the y-pass sums on the FP64 pipe instead of two F2F per tap — parity, then bench c1
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_reference.py tests/test_tiles.py -m gpu -q --maxfail=8 -k "not many_seeds" > gpurun_out/s20_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s20_pytest.txt; tail -6 gpurun_out/s20_pytest.txt | cut -c1-250
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/s20_bench.json 2> gpurun_out/s20_bench.err; echo "bench rc $?"; tail -3 gpurun_out/s20_bench.err
python - <<'PY'
import json
r=json.loads([l for l in open('gpurun_out/s20_bench.json') if l.startswith('{')][-1])
print(r['value'], r['e2e']['value'], r['stage_ms_serial'])
PY
