python -m pytest tests -m gpu -x -q 2>&1 | tail -3
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,1),'frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'], 'e2e %.3g'%d['e2e']['value'])"; }
for t in 0 32 64 128; do PGA_HOOK_THREADS=$t python bench.py --workload test1 --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "test1 hook_threads=$t"; done
python bench.py --workload tsp --steps 2 --warmup 3 --gens 5 --no-cpu-baseline 2>/dev/null | show "tsp"
