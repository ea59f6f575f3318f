python -m pytest tests/test_gpu_nccl.py tests/test_gpu_islands.py tests/test_zz_no_alloc.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
for r in 0 8; do
PGA_RESERVE_SMS=$r python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29800+r)) bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('reserve=$r N=2 value %.4g evals/s  ms/step %.3f' % (d['value'], d['ms_per_step']))"
done
python bench.py --steps 10 --warmup 3 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('N=1 value %.4g evals/s  ms/step %.3f' % (d['value'], d['ms_per_step']))"
python bench.py --steps 5 --warmup 3 --islands 2 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('N=1 2 islands m=10 value %.4g evals/s  ms/step %.3f' % (d['value'], d['ms_per_step']))"
