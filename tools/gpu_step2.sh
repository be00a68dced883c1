#!/bin/bash
# arm B of the k_convolve arithmetic comparison (packed FP32 products): time and worst pixel at full size
mkdir -p gpurun_out
ZB_LIB=$PWD/build/libzb200_fp32.so timeout 900 python tools/conv_ab.py fp32 c1 c1b c4 > gpurun_out/s2_conv_fp32.txt 2>&1
echo "rc $?" >> gpurun_out/s2_conv_fp32.txt
cat gpurun_out/s2_conv_fp32.txt
