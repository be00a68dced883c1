#!/bin/bash
# k_build_matrix_main / _edges (bulk of the normal matrix with shared w*w products) and four chains per thread in
# k_lu_trail: parity of the small cases, serial stage times against step 24's base arm, c1b and c1 bench lines
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "small_cases_host or full_size" > gpurun_out/s25_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s25_pytest.txt; tail -3 gpurun_out/s25_pytest.txt | cut -c1-200
: > gpurun_out/s25_fit_ab.txt
timeout 300 python tools/fit_ab.py new c1 c1b c4 >> gpurun_out/s25_fit_ab.txt 2>&1; echo "new rc $?" >> gpurun_out/s25_fit_ab.txt
ZB_BM_LEGACY=1 timeout 300 python tools/fit_ab.py legacy_bm c1b >> gpurun_out/s25_fit_ab.txt 2>&1; echo "legacy rc $?" >> gpurun_out/s25_fit_ab.txt
cut -c1-420 gpurun_out/s25_fit_ab.txt
for w in c1b c1; do
  timeout 600 python bench.py --workload $w --steps 4 --warmup 3 --no-cpu > gpurun_out/s25_bench_$w.json 2> gpurun_out/s25_bench_$w.err; echo "bench $w rc $?"
done
python - <<'PY'
import json
for w in ("c1b", "c1"):
    try:
        r = json.loads([l for l in open('gpurun_out/s25_bench_%s.json' % w) if l.startswith('{')][-1])
        print(w, r['value'], r['e2e']['value'], r.get('stage_ms_serial'))
    except Exception as e:
        print(w, "no line:", e)
PY
