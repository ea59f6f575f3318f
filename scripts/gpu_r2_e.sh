# round 2, GPU call E: draws-first variants of the register-tiled kernel (PGA_DF=<tiles per batch><blocks per SM>)
mkdir -p gpurun_out
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2),'frac', round(d['roofline']['frac'],3), 'evals/s %.4g'%d['value'])"; }
for df in 0 82 42 43 23 24; do
  PGA_DF=$df timeout 120 python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline --no-c5 2>/dev/null | show "onemax 2^20 df=$df"
done 2>&1 | tee gpurun_out/r2e_df.log
for df in 82 43; do
PGA_DF=$df timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -x -q -k "onemax or golden" 2>&1 | tail -2
for sz in 262144 4194304; do
  PGA_DF=$df timeout 120 python bench.py --size $sz --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax size=$sz df=$df"
done 2>&1 | tee -a gpurun_out/r2e_df.log
done
