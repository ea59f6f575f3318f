python -m pytest tests -m gpu -x -q 2>&1 | tail -3
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,1),'frac', round(d['roofline']['frac'],3), 'evals/s %.3g'%d['value'])"; }
for nc in 1 0; do
PGA_NO_CLASS_TABLE=$nc python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax no_class_table=$nc"
PGA_NO_CLASS_TABLE=$nc python bench.py --workload rastrigin --steps 3 --warmup 3 --gens 20 --no-cpu-baseline 2>/dev/null | show "rastrigin no_class_table=$nc"
PGA_NO_CLASS_TABLE=$nc python bench.py --steps 3 --warmup 3 --gens 20 --size 4194304 --no-cpu-baseline 2>/dev/null | show "onemax 2^22 no_class_table=$nc"
done
