# round 2, 2-GPU call: NCCL islands parity (incl. early stop over ranks), the bench line at N=2 with c5 and islands_parity
mkdir -p gpurun_out
nvidia-smi -L
timeout 600 python -m pytest tests/test_gpu_nccl.py -x -q 2>&1 | tail -5
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'evals/s %.4g'%d['value'], 'ms/step', round(d['ms_per_step'],3), 'e2e %.4g'%d['e2e']['value'], 'c5', d.get('c5',{}).get('value'), d.get('c5',{}).get('ms_per_step'), 'parity', d.get('islands_parity'))"; }
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2h_n1.json 2>gpurun_out/r2h_n1.err; show N=1 < gpurun_out/r2h_n1.json
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29541 bench.py --gpus 2 --steps 10 --warmup 3 > gpurun_out/r2h_n2.json 2>gpurun_out/r2h_n2.err; show N=2 < gpurun_out/r2h_n2.json; tail -3 gpurun_out/r2h_n2.err
for r in 4 8 16; do
PGA_RESERVE_SMS=$r timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 2955$r bench.py --gpus 2 --steps 10 --warmup 3 --no-c5 2>/dev/null | show "N=2 reserve=$r"
done
timeout 600 python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port 29561 bench.py --gpus 2 --impl reference --steps 2 --warmup 1 2>/dev/null | cut -c1-600
