#!/bin/bash
# launch list of one context running c1 subtractions serially (cold-cache, serialised: shares, not absolutes)
mkdir -p gpurun_out
CMD="python bench.py --steps 1 --warmup 1 --batch 2 --workers 1 --no-cpu --no-clocks"
timeout 600 $CMD > gpurun_out/s10_plain.json 2> gpurun_out/s10_plain.err; echo "plain rc $?"
timeout 900 ncu --metrics gpu__time_duration.sum,launch__grid_size,launch__block_size --clock-control none -c 2500 --csv --log-file gpurun_out/s10_launches.csv $CMD > gpurun_out/s10_ncu.log 2>&1; echo "ncu rc $?"
tail -2 gpurun_out/s10_ncu.log | cut -c1-300
