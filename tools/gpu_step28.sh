#!/bin/bash
# launch list of c1b on the final tree; bench lines of c1_dir_b (hotpants' default direction choice) and c4
mkdir -p gpurun_out
CMD="python bench.py --workload c1b --steps 1 --warmup 1 --batch 2 --workers 1 --no-cpu --no-clocks"
timeout 900 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size --clock-control none -c 6000 --csv --log-file gpurun_out/s28_launches_c1b.csv $CMD > gpurun_out/s28_ncu.log 2>&1; echo "ncu rc $?"
python tools/launch_summary.py gpurun_out/s28_launches_c1b.csv > gpurun_out/s28_launches_c1b_summary.txt 2>&1; head -22 gpurun_out/s28_launches_c1b_summary.txt | cut -c1-160
for w in c1_dir_b c4; do
  timeout 600 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu > gpurun_out/s28_bench_$w.json 2> gpurun_out/s28_bench_$w.err; echo "bench $w rc $?"
done
python - <<'PY'
import json
for w in ("c1_dir_b", "c4"):
    try:
        r = json.loads([l for l in open('gpurun_out/s28_bench_%s.json' % w) if l.startswith('{')][-1])
        print(w, r['value'], r['e2e']['value'], r.get('stage_ms_serial'))
    except Exception as e:
        print(w, "no line:", e); print(open('gpurun_out/s28_bench_%s.err' % w).read()[-1500:])
PY
