#!/bin/bash
# whole GPU suite + c1 bench after: conv mask work split over bands (prefix sums kept), two shell passes in the region statistics, pipelined back-substitution
mkdir -p gpurun_out
timeout 3000 python -m pytest tests -m gpu -q --maxfail=12 > gpurun_out/s13_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s13_pytest.txt; tail -12 gpurun_out/s13_pytest.txt | cut -c1-220
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/s13_bench.json 2> gpurun_out/s13_bench.err; echo "bench rc $?"; tail -3 gpurun_out/s13_bench.err
python - <<'PY'
import json
r=json.loads([l for l in open('gpurun_out/s13_bench.json') if l.startswith('{')][-1])
print(r['value'], r['e2e']['value'], r['stage_ms_serial'])
PY
