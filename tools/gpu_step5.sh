#!/bin/bash
# timing of k_convolve at full size + the whole GPU test suite + a short bench
mkdir -p gpurun_out
NO_ORACLE=1 timeout 600 python tools/conv_ab.py hybrid c1 c1b c4 > gpurun_out/s5_conv.txt 2>&1; echo "rc $?" >> gpurun_out/s5_conv.txt; cut -c1-110 gpurun_out/s5_conv.txt
timeout 2400 python -m pytest tests -m gpu -x -q > gpurun_out/s5_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s5_pytest.txt; tail -15 gpurun_out/s5_pytest.txt
timeout 900 python bench.py --steps 3 --warmup 3 > gpurun_out/s5_bench.json 2> gpurun_out/s5_bench.err; echo "bench rc $?"; cut -c1-3000 gpurun_out/s5_bench.json; tail -5 gpurun_out/s5_bench.err
