show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2))"; }
for sm in 0 1024 65536 131072 163840 200000; do
PGA_DUMMY_SMEM=$sm python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline --no-c5 2>/dev/null | show "onemax dummy_smem=$sm"
done
