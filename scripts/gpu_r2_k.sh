# new cooperative select (1 block/SM, 3 barriers, histogram from the generation kernel): tests, then timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2k_pytest.log 2>&1; tail -3 gpurun_out/r2k_pytest.log
python scripts/dbg_select.py 20000; python scripts/dbg_select.py 1048576; python scripts/dbg_select.py 4194304
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2),'frac', round(d['roofline']['frac'],3), 'evals/s %.4g'%d['value'], 'ms/step', round(d['ms_per_step'],3), 'c5', d.get('c5',{}).get('ms_per_step'))"; }
python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax N=1"
python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline --no-c5 --islands 2 2>/dev/null | show "onemax islands=2 one GPU"
PGA_TRACE=1 python bench.py --steps 2 --warmup 3 --gens 100 --no-cpu-baseline --no-c5 --islands 2 2>&1 | grep pga-trace | tail -12
