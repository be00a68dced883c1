timeout 130 python -m pytest tests -x -q -m gpu 2>&1 | tail -2
for r in 0 1; do
ZB_QA_ROT=$r timeout 60 python bench.py --steps 3 --no-cpu 2>/dev/null | grep '^{' | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('rot=$r', round(j['value'],1), round(j['e2e']['value'],1), j['stage_ms_serial']['region_stats'])"
done
