python -m pytest tests -m gpu -x -q 2>&1 | tail -3
B="python bench.py --steps 3 --warmup 3 --gens 100 --no-cpu-baseline"
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,1),'frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'], 'e2e %.3g'%d['e2e']['value'])"; }
for sw in 4 8; do PGA_STAGED_WARPS=$sw $B 2>/dev/null | show "onemax staged warps=$sw"; done
for mb in 2 3; do PGA_NO_STAGED=1 PGA_TILED_MIN_BLOCKS=$mb $B 2>/dev/null | show "onemax tiled minb=$mb (auto grid)"; done
B="python bench.py --workload rastrigin --steps 3 --warmup 3 --gens 20 --no-cpu-baseline"
for mb in 2 3; do PGA_TILED_MIN_BLOCKS=$mb $B 2>/dev/null | show "rastrigin tiled minb=$mb (auto grid)"; done
