for flag in "" "--no-pin"; do
taskset -c 0-7 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 --no-cpu $flag 2>/dev/null | grep "^{" | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('8 cores', '$flag', round(j['value'],1), round(j['e2e']['value'],1), j['config']['host_cores'], j['config']['rank_cores'])"
done
python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29511 bench.py --gpus 2 --steps 4 --warmup 3 --no-cpu 2>/dev/null | grep "^{" | python -c "import sys,json; j=json.loads(sys.stdin.read()); print('all cores', round(j['value'],1), round(j['e2e']['value'],1), j['config']['host_cores'], j['config']['rank_cores'])"
