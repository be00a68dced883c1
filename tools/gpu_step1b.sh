#!/bin/bash
mkdir -p gpurun_out
( for a in "512 512 64 64 0 0" "512 512 84 62 0 0" "512 512 84 62 21 42" "512 512 84 62 462 480" "512 512 84 62 0 0 20864" "512 512 64 16 0 0" "3072 3080 84 62 1000 1000"; do
    echo "== $a"; timeout 60 ./build/tma_probe $a; echo "rc $?"; done ) > gpurun_out/s1_tma_probe.txt 2>&1
cat gpurun_out/s1_tma_probe.txt
