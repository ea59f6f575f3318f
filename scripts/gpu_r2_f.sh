# round 2, GPU call F: cleaned-up library (one block per SM, 5-launch select beside a generation),
# first-parent-ordered generation for long rows (PGA_SORTED), one-GPU island timing
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest.log 2>&1; tail -3 gpurun_out/r2f_pytest.log
PGA_SORTED=1 timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2f_pytest_sorted.log 2>&1; tail -3 gpurun_out/r2f_pytest_sorted.log
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2),'frac', round(d['roofline']['frac'],3), 'evals/s %.4g'%d['value'], 'ms/step', round(d['ms_per_step'],3))"; }
for srt in 0 1 2; do
  PGA_SORTED=$srt timeout 120 python bench.py --workload rastrigin --steps 3 --warmup 3 --gens 30 --no-cpu-baseline 2>/dev/null | show "rastrigin 2^20 sorted=$srt"
done 2>&1 | tee gpurun_out/r2f_sorted.log
python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline --no-c5 2>/dev/null | show "onemax 2^20 one block per SM" | tee -a gpurun_out/r2f_sorted.log
for isl in 1 2; do
  python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline --no-c5 --islands $isl 2>/dev/null | show "onemax 2^20 islands=$isl (one GPU)"
done 2>&1 | tee gpurun_out/r2f_islands.log
