# 2 GPUs: where an exchange's time goes (PGA_TRACE)
PGA_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 3 --warmup 3 --no-c5 2>&1 | grep "pga-trace" | tail -24
PGA_TRACE=1 PGA_NO_COOP=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29542 bench.py --gpus 2 --steps 3 --warmup 3 --no-c5 2>&1 | grep "pga-trace" | tail -12
