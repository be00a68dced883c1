#!/bin/bash
# -c b (check_stamps) on the GPU: the five small cases against oracle + golden, the full-size mixed-direction case against
# the reference binary; then the pending A/B of k_convolve's sleeping waits, the whole GPU suite and a bench line.
mkdir -p gpurun_out
timeout 600 python -m pytest tests/test_gpu_parity.py -m gpu -q -x -k "dir_b" > gpurun_out/s22_dirb.txt 2>&1; echo "dir_b rc $?" >> gpurun_out/s22_dirb.txt; tail -25 gpurun_out/s22_dirb.txt | cut -c1-300
timeout 600 python -m pytest tests/test_gpu_reference.py -m gpu -q -x -s -k "c1_dir_b" > gpurun_out/s22_dirb_full.txt 2>&1; echo "dir_b full rc $?" >> gpurun_out/s22_dirb_full.txt; tail -12 gpurun_out/s22_dirb_full.txt | cut -c1-300
: > gpurun_out/s22_ab.txt
NO_ORACLE=1 timeout 200 python tools/conv_ab.py sleep c1 c1b >> gpurun_out/s22_ab.txt 2>&1; echo "sleep rc $?" >> gpurun_out/s22_ab.txt
ZB_LIB=build/libzb200_spin.so NO_ORACLE=1 timeout 200 python tools/conv_ab.py spin c1 c1b >> gpurun_out/s22_ab.txt 2>&1; echo "spin rc $?" >> gpurun_out/s22_ab.txt
cut -c1-200 gpurun_out/s22_ab.txt
timeout 1500 python -m pytest tests -x -q -m gpu > gpurun_out/s22_pytest.txt 2>&1; echo "pytest rc $?" >> gpurun_out/s22_pytest.txt; tail -6 gpurun_out/s22_pytest.txt | cut -c1-250
timeout 600 python bench.py --steps 4 --warmup 3 > gpurun_out/s22_bench.json 2> gpurun_out/s22_bench.err; echo "bench rc $?"
tail -c 1500 gpurun_out/s22_bench.json
