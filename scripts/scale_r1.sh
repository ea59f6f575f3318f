# 1 -> 8 GPU scaling of the headline workload (one 2^20 x 1024-bit OneMax island per GPU,
# ring migration of the top 1 % every 10 generations over NCCL), then BASELINE config 5
# (8 islands x 2^22 spread over the GPUs)
nvidia-smi -L | wc -l
NG=$(nvidia-smi -L | wc -l)
python -m pytest tests/test_gpu_nccl.py -m gpu -x -q 2>&1 | tail -3
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29600+n)) bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/scale_$n.json 2> gpurun_out/scale_$n.err; fi
  python -c "import json; d=json.loads(open('gpurun_out/scale_$n.json').readline()); print('N=$n value %.4g evals/s  ms/step %.3f  e2e %.4g' % (d['value'], d['ms_per_step'], d['e2e']['value']))" || tail -5 gpurun_out/scale_$n.err
done
for n in 1 2 4 8; do
  [ $n -le $NG ] || continue
  K=$((8/n))
  if [ $n -eq 1 ]; then python bench.py --gpus 1 --steps 3 --warmup 3 --gens 20 --size 4194304 --islands $K --no-cpu-baseline > gpurun_out/c5_$n.json 2> gpurun_out/c5_$n.err
  else python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port $((29700+n)) bench.py --gpus $n --steps 3 --warmup 3 --gens 20 --size 4194304 --islands $K > gpurun_out/c5_$n.json 2> gpurun_out/c5_$n.err; fi
  python -c "import json; d=json.loads(open('gpurun_out/c5_$n.json').readline()); print('C5 N=$n islands/GPU=$K value %.4g evals/s  ms/step %.3f' % (d['value'], d['ms_per_step']))" || tail -5 gpurun_out/c5_$n.err
done
