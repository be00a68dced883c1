#!/bin/bash
# k_lu_subst with warp 0 running the chain only: parity of the small cases, serial stage times and solution hashes
mkdir -p gpurun_out
timeout 300 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "small_cases_host" > gpurun_out/s35_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s35_pytest.txt; tail -2 gpurun_out/s35_pytest.txt | cut -c1-200
timeout 200 python tools/fit_ab.py subst2 c1 c1b c4 > gpurun_out/s35_fit_ab.txt 2>&1; echo "rc $?" >> gpurun_out/s35_fit_ab.txt
cut -c1-400 gpurun_out/s35_fit_ab.txt
