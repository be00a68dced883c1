#!/bin/bash
# hybrid k_convolve (image FP64 / variance FP32): correctness on small cases, time + worst pixel at full size, one ncu capture
mkdir -p gpurun_out
timeout 300 python tests/gpu_case.py c1_small zuds_actual_2x2 conv_image gain_model two_gauss_3x3 > gpurun_out/s3_case.txt 2>&1
echo "rc $?" >> gpurun_out/s3_case.txt
tail -12 gpurun_out/s3_case.txt | cut -c1-300
grep -q "rc 0" gpurun_out/s3_case.txt || exit 0
timeout 900 python tools/conv_ab.py hybrid c1 c1b c4 > gpurun_out/s3_conv_hybrid.txt 2>&1
echo "rc $?" >> gpurun_out/s3_conv_hybrid.txt
cat gpurun_out/s3_conv_hybrid.txt
NO_ORACLE=1 timeout 300 python tools/conv_ab.py hybrid c1 > gpurun_out/s3_plain.log 2>&1 && \
NO_ORACLE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_convolve -c 1 -o gpurun_out/s3_prof_conv python tools/conv_ab.py hybrid c1 > gpurun_out/s3_ncu.log 2>&1
echo "ncu rc $?"; tail -3 gpurun_out/s3_ncu.log
