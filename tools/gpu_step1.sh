#!/bin/bash
# first B200 contact of the rebuilt k_convolve: microbenchmark, one small case under a short timeout, then the suite
mkdir -p gpurun_out
nvidia-smi --query-gpu=name,clocks.sm,clocks.max.sm,power.draw --format=csv > gpurun_out/s1_smi.txt 2>&1
./build/fp32_peak > gpurun_out/s1_fp32_peak.txt 2>&1
echo "== one small case" > gpurun_out/s1_case.txt
timeout 180 python tests/gpu_case.py c1_small zuds_actual_2x2 conv_image >> gpurun_out/s1_case.txt 2>&1
echo "rc $?" >> gpurun_out/s1_case.txt
tail -30 gpurun_out/s1_case.txt
if grep -q "rc 0" gpurun_out/s1_case.txt; then
  timeout 1500 python -m pytest tests -m gpu -x -q -k "not full_size_against and not many_seeds and not nccl" > gpurun_out/s1_pytest.txt 2>&1
  echo "pytest rc $?" >> gpurun_out/s1_pytest.txt
  tail -25 gpurun_out/s1_pytest.txt
  timeout 600 python bench.py --steps 3 --warmup 3 --no-cpu > gpurun_out/s1_bench.txt 2> gpurun_out/s1_bench.err
  echo "bench rc $?"; cat gpurun_out/s1_bench.txt | cut -c1-1500
fi
