#!/bin/bash
# the whole GPU test suite (all failures listed) + conv timing
mkdir -p gpurun_out
timeout 3000 python -m pytest tests -m gpu -q --maxfail=12 > gpurun_out/s5_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s5_pytest.txt; tail -40 gpurun_out/s5_pytest.txt | cut -c1-220
NO_ORACLE=1 timeout 600 python tools/conv_ab.py hybrid c1 c1b c4 > gpurun_out/s5_conv.txt 2>&1; echo "rc $?" >> gpurun_out/s5_conv.txt; cut -c1-110 gpurun_out/s5_conv.txt
