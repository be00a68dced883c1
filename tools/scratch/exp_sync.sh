run() { # mode workers
  ZB_SYNC=$1 taskset -c 0-3 python bench.py --steps 3 --no-cpu --workers $2 2>/dev/null | python -c "import sys,json; j=json.loads([l for l in sys.stdin if l.startswith('{')][-1]); print('$1', 'W=$2', round(j['value'],1), round(j['e2e']['value'],1))"
}
run spin 4; run spin 8; run yield 8; run yield 12; run yield 24; run block 12; run block 24
for d in 0 1; do ZB_QA_DYN=$d python tools/exp_dma.py 1 2>&1 | grep -i quiet; done
