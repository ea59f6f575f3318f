python -m pytest tests/test_gpu_nccl.py tests/test_gpu_islands.py tests/test_zz_no_alloc.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
PGA_NO_COOP=1 python -m pytest tests/test_gpu_islands.py -m gpu -x -q 2>&1 | tail -2
for nc in 0 1; do
PGA_NO_COOP=$nc python -m torch.distributed.run --nnodes=1 --nproc-per-node 2 --master-addr 127.0.0.1 --master-port $((29800+nc)) bench.py --gpus 2 --steps 10 --warmup 3 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('no_coop=$nc N=2 value %.4g evals/s  ms/step %.3f' % (d['value'], d['ms_per_step']))"
done
python bench.py --steps 5 --warmup 3 --islands 2 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('N=1 2 islands m=10 value %.4g evals/s  ms/step %.3f' % (d['value'], d['ms_per_step']))"
