#!/bin/bash
# conv changes (mask work split over bands, no prefix sums, V rows by mbarrier, convert unrolled): parity, bench, launch list
mkdir -p gpurun_out
timeout 1500 python -m pytest tests/test_gpu_parity.py tests/test_gpu_reference.py tests/test_tiles.py -m gpu -q --maxfail=8 -k "not many_seeds" > gpurun_out/s11_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s11_pytest.txt; tail -8 gpurun_out/s11_pytest.txt | cut -c1-250
timeout 600 python bench.py --steps 4 --warmup 3 --no-cpu > gpurun_out/s11_bench.json 2> gpurun_out/s11_bench.err; echo "bench rc $?"; tail -3 gpurun_out/s11_bench.err
python - <<'PY'
import json
r=json.loads([l for l in open('gpurun_out/s11_bench.json') if l.startswith('{')][-1])
print(r['value'], r['e2e']['value'], r['stage_ms_serial'])
PY
CMD="python bench.py --steps 1 --warmup 1 --batch 2 --workers 1 --no-cpu --no-clocks"
timeout 600 $CMD > gpurun_out/s11_plain.json 2> gpurun_out/s11_plain.err; echo "plain rc $?"
timeout 900 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size --clock-control none -c 2500 --csv --log-file gpurun_out/s11_launches.csv $CMD > gpurun_out/s11_ncu.log 2>&1; echo "ncu rc $?"
tail -2 gpurun_out/s11_ncu.log | cut -c1-300
