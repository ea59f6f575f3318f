# 2 GPUs: tests, N=1 and N=2 bench lines, trace of an exchange
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2l_pytest.log 2>&1; tail -3 gpurun_out/r2l_pytest.log
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'evals/s %.4g'%d['value'], 'ms/step', round(d['ms_per_step'],3), 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2), 'e2e %.4g'%d['e2e']['value'], 'c5', d.get('c5',{}).get('value'), d.get('c5',{}).get('ms_per_step'), 'parity', (d.get('islands_parity') or {}).get('ok'))"; }
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2l_n1.json 2>gpurun_out/r2l_n1.err; show N=1 < gpurun_out/r2l_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2l_n2.json 2>gpurun_out/r2l_n2.err; show N=2 < gpurun_out/r2l_n2.json
PGA_RESERVE_SMS=4 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29551 bench.py --gpus 2 --steps 10 --warmup 3 --no-c5 2>/dev/null | show "N=2 reserve=4"
PGA_TRACE=1 timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29543 bench.py --gpus 2 --steps 3 --warmup 3 --no-c5 2>&1 | grep "pga-trace" | head -12
