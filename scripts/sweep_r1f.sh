python -m pytest tests/test_gpu_islands.py tests/test_gpu_golden.py -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 5 --warmup 3 --no-cpu-baseline"
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'ms/step', round(d['ms_per_step'],3), 'us/gen-launch', round(d['roofline']['avg_launch_ms']*1000,1), 'evals/s %.3g'%d['value'], 'launches', d['gpu_launches'])"; }
$B --gens 100 2>/dev/null | show "1 island pga_run"
$B --gens 100 --islands 2 --migrate-every 1000 2>/dev/null | show "2 islands no migration"
$B --gens 100 --islands 2 2>/dev/null | show "2 islands ring m=10"
$B --gens 100 --islands 2 --migrate-every 1 2>/dev/null | show "2 islands ring m=1"
$B --gens 20 --size 4194304 --islands 8 2>/dev/null | show "C5 on 1 GPU: 8 x 2^22 ring m=10"
$B --gens 20 --size 4194304 --islands 1 2>/dev/null | show "1 x 2^22 pga_run"
