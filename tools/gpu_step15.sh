#!/bin/bash
# ncu --set full of one warm k_convolve launch at c1 (after the same command ran clean without ncu)
mkdir -p gpurun_out
NO_ORACLE=1 timeout 300 python tools/conv_ab.py final c1 > gpurun_out/s15_plain.log 2>&1; echo "plain rc $?"; cat gpurun_out/s15_plain.log | cut -c1-200
NO_ORACLE=1 timeout 900 ncu --set full --clock-control none --import-source on -k regex:k_convolve -s 2 -c 1 -f -o gpurun_out/s15_prof_conv python tools/conv_ab.py final c1 > gpurun_out/s15_ncu.log 2>&1
echo "ncu rc $?"; tail -3 gpurun_out/s15_ncu.log
