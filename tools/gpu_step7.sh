#!/bin/bash
# 2 GPUs: pair sharding (weak scaling) and the row-tile path (strong scaling) through bench.py as the driver launches it, NCCL tile test
mkdir -p gpurun_out
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29611"
timeout 900 $TR bench.py --gpus 2 --steps 3 --warmup 3 > gpurun_out/s7_bench_n2.json 2> gpurun_out/s7_bench_n2.err; echo "n2 rc $?"; cut -c1-300 gpurun_out/s7_bench_n2.json; tail -3 gpurun_out/s7_bench_n2.err
timeout 900 $TR bench.py --gpus 2 --workload c5 --steps 3 --warmup 2 --check > gpurun_out/s7_bench_c5_n2.json 2> gpurun_out/s7_bench_c5_n2.err; echo "c5 n2 rc $?"; cut -c1-300 gpurun_out/s7_bench_c5_n2.json; tail -3 gpurun_out/s7_bench_c5_n2.err
timeout 900 python -m pytest tests/test_gpu_reference.py -q -k nccl > gpurun_out/s7_pytest_nccl.txt 2>&1; tail -5 gpurun_out/s7_pytest_nccl.txt
