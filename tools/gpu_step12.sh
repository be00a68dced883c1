#!/bin/bash
# A/B of the k_convolve changes (timing only): the library as built, mask work on band 0 only, everything as before
mkdir -p gpurun_out; : > gpurun_out/s12_ab.txt
NO_ORACLE=1 timeout 300 python tools/conv_ab.py new c1 c1b c4 >> gpurun_out/s12_ab.txt 2>&1; echo "new rc $?" >> gpurun_out/s12_ab.txt
for v in band0 allold; do
  ZB_LIB=build/libzb200_$v.so NO_ORACLE=1 timeout 300 python tools/conv_ab.py $v c1 >> gpurun_out/s12_ab.txt 2>&1; echo "$v rc $?" >> gpurun_out/s12_ab.txt
done
cut -c1-160 gpurun_out/s12_ab.txt
