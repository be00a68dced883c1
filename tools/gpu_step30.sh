#!/bin/bash
# whole GPU suite on the tree with the staged normal-matrix kernels and the per-region clip CTA
mkdir -p gpurun_out
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/s30_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s30_pytest.txt; tail -4 gpurun_out/s30_pytest.txt | cut -c1-250
