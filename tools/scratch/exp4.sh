for w in 32 48; do
python bench.py --steps 4 --no-cpu --workers $w 2>/dev/null | grep '^{' | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('W=$w', round(j['value'],1), round(j['e2e']['value'],1), j['config']['pairs_per_step_per_gpu'])"
done
python -m pytest tests -x -q -m gpu 2>&1 | tail -2
