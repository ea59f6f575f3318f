# round 2, GPU call A: full GPU test suite (with the reference pin), the round-2
# microbenchmarks, the OneMax kernel at L2-resident sizes, L2 persistence windows,
# and one full ncu capture of the headline kernel
set -x
mkdir -p gpurun_out
nvidia-smi -L
timeout 1500 python -m pytest tests -m gpu -x -q > gpurun_out/r2a_pytest.log 2>&1; tail -15 gpurun_out/r2a_pytest.log
timeout 300 scripts/microbench_r2.bin ABCDE > gpurun_out/r2a_microbench.log 2>&1; cat gpurun_out/r2a_microbench.log
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2),'frac', round(d['roofline']['frac'],3), 'evals/s %.4g'%d['value'], 'ns/1k children', round(d['roofline']['avg_launch_ms']*1e6/ $2 *1000, 2))"; }
for sz in 131072 262144 524288 1048576 2097152 4194304; do
  python bench.py --size $sz --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax size=$sz" $sz
done 2>&1 | tee gpurun_out/r2a_sizes.log
for mode in 1 2 5 6; do
  PGA_L2_PERSIST=$mode python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax 2^20 l2_persist=$mode" 1048576
done 2>&1 | tee gpurun_out/r2a_persist.log
for sz in 262144 1048576; do
  python bench.py --workload rastrigin --size $sz --steps 3 --warmup 3 --gens 20 --no-cpu-baseline 2>/dev/null | show "rastrigin size=$sz" $sz
done 2>&1 | tee -a gpurun_out/r2a_sizes.log
CMD="python bench.py --steps 1 --warmup 3 --gens 5 --no-cpu-baseline"
$CMD > gpurun_out/r2a_plain_full.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:generation_fused_tiled -s 12 -c 2 -o gpurun_out/prof_onemax_r2a $CMD > gpurun_out/r2a_ncu_full.log 2>&1
tail -3 gpurun_out/r2a_ncu_full.log
