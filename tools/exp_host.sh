#!/bin/bash
# Host-side experiments behind profiles/r01_host_threads_s18.md (run on the GPU box: gpurun -- 'bash tools/exp_host.sh sync').
# Every block prints "label value e2e" lines (quadrant-pairs/s, bench.py --no-cpu); compare within one call only.
line() { grep '^{' | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('$1', round(j['value'],1), round(j['e2e']['value'],1))"; }
case "$1" in
sync)      # how a worker thread waits for its stream, with 4 cores for the whole process
  for cfg in "spin 4" "spin 8" "yield 8" "yield 12" "yield 24" "block 12" "block 24"; do
    set -- $cfg
    ZB_SYNC=$1 taskset -c 0-3 python bench.py --steps 3 --no-cpu --workers $2 2>/dev/null | line "4cores $1 W=$2"
  done ;;
queues)    # CUDA hardware work queues vs contexts
  for c in 8 32; do for w in 24 32 48; do
    CUDA_DEVICE_MAX_CONNECTIONS=$c python bench.py --steps 4 --no-cpu --workers $w 2>/dev/null | line "queues=$c W=$w"
  done; done ;;
pin)       # two ranks on 8 cores, bound to 4 each or not (needs gpurun --gpus 2)
  for flag in "" "--no-pin"; do
    taskset -c 0-7 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 \
      bench.py --gpus 2 --steps 4 --warmup 3 --no-cpu $flag 2>/dev/null | line "2gpus 8cores ${flag:-pinned}"
  done ;;
qa)        # k_qa_pass generic form unrolled (0) or rotated (1)
  for r in 0 1; do ZB_QA_ROT=$r python bench.py --steps 3 --no-cpu 2>/dev/null | line "ZB_QA_ROT=$r"; done ;;
*) echo "usage: $0 sync|queues|pin|qa"; exit 2 ;;
esac
