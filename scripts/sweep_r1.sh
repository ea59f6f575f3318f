python -m pytest tests -m gpu -x -q 2>&1 | tail -5
B="python bench.py --steps 3 --warmup 3 --gens 100 --no-cpu-baseline"
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'ms/gen', round(d['roofline']['avg_launch_ms']*1000,1),'us frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'], 'e2e %.3g'%d['e2e']['value'])"; }
for ld in 0 1 2; do for st in 0 1 2; do PGA_L2_LOAD=$ld PGA_L2_STORE=$st $B 2>/dev/null | show "onemax ld=$ld st=$st"; done; done
for bps in 2 4 8 32; do PGA_FAST_BLOCKS_PER_SM=$bps $B 2>/dev/null | show "onemax bps=$bps"; done
PGA_FORCE_GENERIC=1 $B 2>/dev/null | show "onemax generic"
B="python bench.py --workload rastrigin --steps 3 --warmup 3 --gens 20 --no-cpu-baseline"
for ld in 0 1; do for st in 0 2; do PGA_L2_LOAD=$ld PGA_L2_STORE=$st $B 2>/dev/null | show "rastrigin ld=$ld st=$st"; done; done
