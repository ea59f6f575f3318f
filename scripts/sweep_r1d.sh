B="python bench.py --steps 3 --warmup 3 --gens 100 --no-cpu-baseline"
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,1),'frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'])"; }
for v in 3,1,4,2 3,1,4,3 3,1,2,3 3,1,2,4 2,2,2,2 2,2,2,3 2,2,1,3 2,2,1,4 2,2,4,2 1,4,1,2 1,4,1,3 1,4,1,4 1,4,2,2 0,8,1,2 0,8,1,3; do PGA_NO_STAGED=1 PGA_TILED_VARIANT=$v $B 2>/dev/null | show "onemax variant LOGG,NB,TB,MINB=$v"; done
$B 2>/dev/null | show "onemax staged (default)"
python -m pytest tests -m gpu -x -q 2>&1 | tail -3
