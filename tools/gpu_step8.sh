#!/bin/bash
# 8 GPUs: end-to-end scaling of the pair-sharded path; sweep of the host-side knobs that could limit it
mkdir -p gpurun_out
nproc > gpurun_out/s8_host.txt; lscpu | grep -i "model name\|socket\|numa\|^cpu(s)" >> gpurun_out/s8_host.txt
nvidia-smi topo -m >> gpurun_out/s8_host.txt 2>&1
free -g | head -2 >> gpurun_out/s8_host.txt
TR="python -m torch.distributed.run --nnodes=1 --nproc-per-node 8 --master-addr 127.0.0.1 --master-port 29611"
run() { tag=$1; shift; timeout 600 env "$@" $TR bench.py --gpus 8 --steps 3 --warmup 3 --no-cpu $EXTRA > gpurun_out/s8_$tag.json 2> gpurun_out/s8_$tag.err; echo "$tag rc $?"; grep '^{' gpurun_out/s8_$tag.json | python -c "
import json,sys
r=json.loads(sys.stdin.readline()); print(round(r['value'],1), round(r['e2e']['value'],1), r['e2e']['host_link'], r['run']['rank_cores'], r['clocks']['sm_mhz'])"; }
EXTRA="" run default ZB_DUMMY=1
EXTRA="" run block ZB_SYNC=block
EXTRA="--workers 16" run w16 ZB_DUMMY=1
EXTRA="--workers 8" run w8 ZB_SYNC=block
