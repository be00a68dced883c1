#!/bin/bash
# the no-allocation entry point: its test alone in a fresh process, then the whole GPU suite (it is collected last)
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_zz_no_alloc.py -m gpu -q > gpurun_out/s19_alone.txt 2>&1; echo "alone rc $?" >> gpurun_out/s19_alone.txt; tail -15 gpurun_out/s19_alone.txt | cut -c1-250
timeout 3000 python -m pytest tests -x -q -m gpu > gpurun_out/s19_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s19_pytest.txt; tail -8 gpurun_out/s19_pytest.txt | cut -c1-250
