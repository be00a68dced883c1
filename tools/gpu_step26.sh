#!/bin/bash
# stamp search with a 16-bit mask copy (four CTAs per SM): whole GPU suite, c1 / c1b bench lines
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/s26_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s26_pytest.txt; tail -4 gpurun_out/s26_pytest.txt | cut -c1-250
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/s26_bench_c1.json 2> gpurun_out/s26_bench_c1.err; echo "bench c1 rc $?"
timeout 600 python bench.py --workload c1b --steps 4 --warmup 3 --no-cpu > gpurun_out/s26_bench_c1b.json 2> gpurun_out/s26_bench_c1b.err; echo "bench c1b rc $?"
python - <<'PY'
import json
for w in ("c1", "c1b"):
    try:
        r = json.loads([l for l in open('gpurun_out/s26_bench_%s.json' % w) if l.startswith('{')][-1])
        print(w, r['value'], r['e2e']['value'], r.get('stage_ms_serial'), r['roofline']['frac'])
    except Exception as e:
        print(w, "no line:", e)
PY
