#!/bin/bash
# final tree: smoke(), the whole GPU suite (every test, no -x), the default bench line
mkdir -p gpurun_out
timeout 300 python -c "import __graft_entry__ as g; g.smoke(); print('smoke ok')" > gpurun_out/s32_smoke.txt 2>&1; echo "smoke rc $?"; tail -2 gpurun_out/s32_smoke.txt | cut -c1-200
timeout 1500 python -m pytest tests -q -m gpu > gpurun_out/s32_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s32_pytest.txt; tail -6 gpurun_out/s32_pytest.txt | cut -c1-250
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/s32_bench_c1.json 2> gpurun_out/s32_bench_c1.err; echo "bench c1 rc $?"
python - <<'PY'
import json
r = json.loads([l for l in open('gpurun_out/s32_bench_c1.json') if l.startswith('{')][-1])
print(r['value'], r['e2e']['value'], r['roofline']['frac'], r.get('e2e_api'), r['cpu_baseline']['value'])
PY
