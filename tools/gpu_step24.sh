#!/bin/bash
# U12 solve by column sweeps (variant build): parity + A/B of the serial fit loop; bench lines of c1b / c4 / c5 on the
# final tree; launch list of c1b
mkdir -p gpurun_out
ZB_LIB=build/libzb200_sweep.so timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "small_cases_host" > gpurun_out/s24_sweep_pytest.txt 2>&1; echo "sweep pytest rc $?" >> gpurun_out/s24_sweep_pytest.txt; tail -3 gpurun_out/s24_sweep_pytest.txt | cut -c1-200
: > gpurun_out/s24_fit_ab.txt
timeout 300 python tools/fit_ab.py base c1 c1b c4 >> gpurun_out/s24_fit_ab.txt 2>&1; echo "base rc $?" >> gpurun_out/s24_fit_ab.txt
ZB_LIB=build/libzb200_sweep.so timeout 300 python tools/fit_ab.py sweep c1 c1b c4 >> gpurun_out/s24_fit_ab.txt 2>&1; echo "sweep rc $?" >> gpurun_out/s24_fit_ab.txt
cut -c1-420 gpurun_out/s24_fit_ab.txt
for w in c1b c4; do
  timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu > gpurun_out/s24_bench_$w.json 2> gpurun_out/s24_bench_$w.err; echo "bench $w rc $?"
done
timeout 600 python bench.py --workload c5 --steps 6 --warmup 3 > gpurun_out/s24_bench_c5.json 2> gpurun_out/s24_bench_c5.err; echo "bench c5 rc $?"
python - <<'PY'
import json
for w in ("c1b", "c4", "c5"):
    try:
        r = json.loads([l for l in open('gpurun_out/s24_bench_%s.json' % w) if l.startswith('{')][-1])
        print(w, r['value'], r['e2e']['value'], r.get('stage_ms_serial'), (r.get('roofline') or {}).get('frac'))
    except Exception as e:
        print(w, "no line:", e)
PY
CMD="python bench.py --workload c1b --steps 1 --warmup 1 --batch 2 --workers 1 --no-cpu --no-clocks"
timeout 900 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size --clock-control none -c 6000 --csv --log-file gpurun_out/s24_launches_c1b.csv $CMD > gpurun_out/s24_ncu.log 2>&1; echo "ncu rc $?"
python tools/launch_summary.py gpurun_out/s24_launches_c1b.csv > gpurun_out/s24_launches_c1b_summary.txt 2>&1; head -24 gpurun_out/s24_launches_c1b_summary.txt | cut -c1-160
