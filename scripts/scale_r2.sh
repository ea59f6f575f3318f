# round 2: 1/2/4/8-GPU scaling of the bench line (with c5 and islands_parity) on one box, plus the NCCL parity test
mkdir -p gpurun_out
N=$(nvidia-smi -L | wc -l)
timeout 900 python -m pytest tests/test_gpu_nccl.py -x -q 2>&1 | tail -2
timeout 300 python bench.py --steps 10 --warmup 3 --no-cpu-baseline > gpurun_out/r2_scale_1.json 2>/dev/null
for n in 2 4 8; do
  [ $n -le $N ] || continue
  timeout 900 python -m torch.distributed.run --nnodes=1 --nproc-per-node $n --master-addr 127.0.0.1 --master-port 2960$n bench.py --gpus $n --steps 10 --warmup 3 > gpurun_out/r2_scale_$n.json 2>gpurun_out/r2_scale_$n.err
done
python - <<'PY'
import json,glob
base=None
for n in (1,2,4,8):
    try: d=json.load(open(f'gpurun_out/r2_scale_{n}.json'))
    except Exception as e: print(n,'missing'); continue
    if n==1: base=d
    c5=d.get('c5',{})
    print(f"N={n} evals/s {d['value']:.4g} ms/step {d['ms_per_step']:.3f} x{d['value']/base['value']:.2f} e2e {d['e2e']['value']:.4g} x{d['e2e']['value']/base['e2e']['value']:.2f} | c5 {c5.get('value',0):.4g} ms/step {c5.get('ms_per_step',0):.2f} x{c5.get('value',0)/base['c5']['value']:.2f} | parity {(d.get('islands_parity') or {}).get('ok')}")
PY
