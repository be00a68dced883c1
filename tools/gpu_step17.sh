#!/bin/bash
# contexts per GPU at the fit-dominated workloads (c4: 1 region, 32x32 stamps; c1b: ko 4): is 32 enough?
mkdir -p gpurun_out; : > gpurun_out/s17_workers.txt
for wl in c4 c1b; do for w in 32 64 96; do
  timeout 600 python bench.py --workload $wl --workers $w --batch $((2*w)) --steps 2 --warmup 1 --no-cpu --no-clocks > gpurun_out/s17_tmp.json 2> gpurun_out/s17_tmp.err; rc=$?
  python - "$wl" "$w" "$rc" >> gpurun_out/s17_workers.txt <<'PY'
import json,sys
try:
    r=json.loads([l for l in open('gpurun_out/s17_tmp.json') if l.startswith('{')][-1])
    print(sys.argv[1], 'workers', sys.argv[2], 'value %.1f e2e %.1f'%(r['value'], r['e2e']['value']), 'fit_loop %.1f ms'%r['stage_ms_serial']['fit_loop'])
except Exception as e:
    print(sys.argv[1], sys.argv[2], 'rc', sys.argv[3], 'failed', e, open('gpurun_out/s17_tmp.err').read()[-300:])
PY
done; done
cat gpurun_out/s17_workers.txt
