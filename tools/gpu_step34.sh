#!/bin/bash
# source-level ncu capture of one k_lu_subst launch at n = 722 (c1b): where the per-row cost of the back-substitution goes
mkdir -p gpurun_out
timeout 400 ncu --set full --import-source on --clock-control none -k regex:k_lu_subst -s 1 -c 1 -o gpurun_out/s34_k_lu_subst -f python tools/one_pair.py c1b > gpurun_out/s34_ncu.log 2>&1; echo "ncu rc $?"; tail -3 gpurun_out/s34_ncu.log | cut -c1-200; ls -la gpurun_out/s34_k_lu_subst.ncu-rep
