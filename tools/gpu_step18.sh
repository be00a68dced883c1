#!/bin/bash
# LU launch sequence as a CUDA graph: parity, then A/B (ZB_LU_GRAPH=0/1) at c4, c1b, c1; where the zuds-level call spends its time
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_tiles.py -m gpu -q --maxfail=8 -k "not full_size" > gpurun_out/s18_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s18_pytest.txt; tail -5 gpurun_out/s18_pytest.txt | cut -c1-200
: > gpurun_out/s18_ab.txt
for wl in c4 c1b c1; do for g in 0 1; do
  ZB_LU_GRAPH=$g timeout 600 python bench.py --workload $wl --batch 64 --steps 2 --warmup 1 --no-cpu --no-clocks > gpurun_out/s18_tmp.json 2> gpurun_out/s18_tmp.err; rc=$?
  python - "$wl" "$g" "$rc" >> gpurun_out/s18_ab.txt <<'PY'
import json,sys
try:
    r=json.loads([l for l in open('gpurun_out/s18_tmp.json') if l.startswith('{')][-1])
    print(sys.argv[1], 'graph', sys.argv[2], 'value %.1f e2e %.1f'%(r['value'], r['e2e']['value']), 'fit_loop %.2f ms'%r['stage_ms_serial']['fit_loop'])
except Exception as e:
    print(sys.argv[1], sys.argv[2], 'rc', sys.argv[3], 'failed', e, open('gpurun_out/s18_tmp.err').read()[-300:])
PY
done; done
cat gpurun_out/s18_ab.txt
timeout 600 python tools/profile_api.py 3 > gpurun_out/s18_api_profile.txt 2>&1; echo "api rc $?"; head -45 gpurun_out/s18_api_profile.txt | cut -c1-150
