B="python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline"
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,1),'frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'])"; }
for st in 0 4 6 7; do PGA_STAGED=$st $B 2>/dev/null | show "onemax staged=$st"; done
PGA_STAGED=7 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
