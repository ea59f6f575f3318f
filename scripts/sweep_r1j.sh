export PGA_PREFETCH=1
python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
unset PGA_PREFETCH
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,1),'frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'])"; }
for pf in 0 1; do
PGA_PREFETCH=$pf python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax prefetch=$pf"
PGA_PREFETCH=$pf python bench.py --workload rastrigin --steps 3 --warmup 3 --gens 20 --no-cpu-baseline 2>/dev/null | show "rastrigin prefetch=$pf"
PGA_PREFETCH=$pf python bench.py --steps 3 --warmup 3 --gens 20 --size 4194304 --no-cpu-baseline 2>/dev/null | show "onemax 2^22 prefetch=$pf"
done
