python bench.py --steps 5 --warmup 3 --islands 2 --migrate-every 100000 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('no migration ms/step %.3f' % (d['ms_per_step']))"
for sk in 0 1 2 4 8 16 9 31 27; do
PGA_DEBUG_SKIP=$sk python bench.py --steps 5 --warmup 3 --islands 2 --no-cpu-baseline 2>/dev/null | python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('skip=$sk ms/step %.3f' % (d['ms_per_step']))"
done
