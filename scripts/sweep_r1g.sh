python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,1),'frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'])"; }
B="python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline"
$B 2>/dev/null | show "onemax pipe"
PGA_NO_PIPE=1 $B 2>/dev/null | show "onemax tiled"
PGA_FAST_BLOCKS_PER_SM=1 $B 2>/dev/null | show "onemax pipe bps=1"
B="python bench.py --workload rastrigin --steps 3 --warmup 3 --gens 20 --no-cpu-baseline"
$B 2>/dev/null | show "rastrigin pipe"
PGA_NO_PIPE=1 $B 2>/dev/null | show "rastrigin tiled"
