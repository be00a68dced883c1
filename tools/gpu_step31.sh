#!/bin/bash
# k_lu_swap_u (permutation by warp 0, pivots preloaded) and k_lu_panel (row update eight columns at a time): parity,
# serial stage times / solution hashes, bench lines
mkdir -p gpurun_out
timeout 900 python -m pytest tests/test_gpu_parity.py tests/test_tiles.py -m gpu -q -x > gpurun_out/s31_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s31_pytest.txt; tail -3 gpurun_out/s31_pytest.txt | cut -c1-200
timeout 300 python tools/fit_ab.py lu2 c1 c1b c4 > gpurun_out/s31_fit_ab.txt 2>&1; echo "rc $?" >> gpurun_out/s31_fit_ab.txt
cut -c1-420 gpurun_out/s31_fit_ab.txt
for w in c4 c1b; do
  timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu > gpurun_out/s31_bench_$w.json 2> gpurun_out/s31_bench_$w.err; echo "bench $w rc $?"
done
python - <<'PY'
import json
for w in ("c4", "c1b"):
    try:
        r = json.loads([l for l in open('gpurun_out/s31_bench_%s.json' % w) if l.startswith('{')][-1])
        print(w, r['value'], r['e2e']['value'], r.get('stage_ms_serial'))
    except Exception as e:
        print(w, "no line:", e); print(open('gpurun_out/s31_bench_%s.err' % w).read()[-1500:])
PY
