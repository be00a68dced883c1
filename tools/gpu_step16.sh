#!/bin/bash
# conflict-free window pitches: timing at c1/c1b/c4, small parity run
mkdir -p gpurun_out
NO_ORACLE=1 timeout 300 python tools/conv_ab.py pitch c1 c1b c4 > gpurun_out/s16_ab.txt 2>&1; echo "rc $?" >> gpurun_out/s16_ab.txt; cut -c1-160 gpurun_out/s16_ab.txt
timeout 900 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "small_cases_host or full_size_quadrant" > gpurun_out/s16_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s16_pytest.txt; tail -4 gpurun_out/s16_pytest.txt | cut -c1-200
