#!/bin/bash
# every bench workload once at N = 1 + the reference arm
mkdir -p gpurun_out
for w in c1b c4; do
  timeout 900 python bench.py --workload $w --steps 3 --warmup 3 --no-cpu > gpurun_out/s6_bench_$w.json 2> gpurun_out/s6_bench_$w.err; echo "$w rc $?"; cut -c1-400 gpurun_out/s6_bench_$w.json; tail -3 gpurun_out/s6_bench_$w.err
done
timeout 900 python bench.py --workload c5 --steps 3 --warmup 3 --check > gpurun_out/s6_bench_c5.json 2> gpurun_out/s6_bench_c5.err; echo "c5 rc $?"; cut -c1-600 gpurun_out/s6_bench_c5.json; tail -3 gpurun_out/s6_bench_c5.err
timeout 900 python bench.py --workload batch256 --npairs 32 --steps 2 --warmup 3 --no-cpu > gpurun_out/s6_bench_batch.json 2> gpurun_out/s6_bench_batch.err; echo "batch rc $?"; cut -c1-400 gpurun_out/s6_bench_batch.json; tail -3 gpurun_out/s6_bench_batch.err
timeout 900 python bench.py --impl reference --steps 1 --warmup 0 > gpurun_out/s6_bench_ref.json 2> gpurun_out/s6_bench_ref.err; echo "ref rc $?"; cut -c1-500 gpurun_out/s6_bench_ref.json
