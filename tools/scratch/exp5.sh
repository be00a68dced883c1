for c in 8 32; do for w in 24 32; do
CUDA_DEVICE_MAX_CONNECTIONS=$c python bench.py --steps 4 --no-cpu --workers $w 2>/dev/null | grep '^{' | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('conn=$c W=$w', round(j['value'],1), round(j['e2e']['value'],1))"
done; done
