#!/bin/bash
# B200 contact of the rebuilt k_convolve: small cases under a short timeout, then the suite and a short bench
mkdir -p gpurun_out
echo "== default build, debug sync" > gpurun_out/s1_case.txt
ZB_DEBUG_SYNC=1 timeout 180 python tests/gpu_case.py c1_small >> gpurun_out/s1_case.txt 2>&1
rc=$?; echo "rc $rc" >> gpurun_out/s1_case.txt
if [ $rc -ne 0 ]; then tail -30 gpurun_out/s1_case.txt; exit 0; fi
timeout 300 python tests/gpu_case.py c1_small zuds_actual_2x2 conv_image gain_model two_gauss_3x3 >> gpurun_out/s1_case.txt 2>&1
echo "rc $?" >> gpurun_out/s1_case.txt
tail -40 gpurun_out/s1_case.txt
timeout 1500 python -m pytest tests -m gpu -x -q -k "not full_size_against and not many_seeds and not nccl" > gpurun_out/s1_pytest.txt 2>&1
echo "pytest rc $?" >> gpurun_out/s1_pytest.txt
tail -25 gpurun_out/s1_pytest.txt
