for sk in 0 3 24 27; do
PGA_DEBUG_SKIP=$sk python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29900+sk)) bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('skip=$sk N=2 ms/step %.3f' % (d['ms_per_step']))"
done
PGA_DEBUG_SKIP=0 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29950 bench.py --gpus 2 --steps 10 --warmup 3 --migrate-every 100000 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('no migration N=2 ms/step %.3f' % (d['ms_per_step']))"
