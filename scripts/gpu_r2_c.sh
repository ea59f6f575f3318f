# round 2, GPU call C: the self-pipelined staged kernel (PGA_WS=8..14 = warps per SM)
set -x
mkdir -p gpurun_out
PGA_WS=13 timeout 300 python -m pytest tests/test_gpu_parity.py tests/test_gpu_golden.py -x -q -k "onemax or golden" > gpurun_out/r2c_pytest_ws.log 2>&1; tail -4 gpurun_out/r2c_pytest_ws.log
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2),'frac', round(d['roofline']['frac'],3), 'evals/s %.4g'%d['value'])"; }
for ws in 0 8 10 12 13 14; do
  PGA_WS=$ws timeout 120 python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline --no-c5 2>/dev/null | show "onemax 2^20 ws=$ws"
done 2>&1 | grep -v "^+" | tee gpurun_out/r2c_ws.log
for sz in 262144 524288 4194304; do
  PGA_WS=13 timeout 120 python bench.py --size $sz --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax size=$sz ws=13"
done 2>&1 | grep -v "^+" | tee -a gpurun_out/r2c_ws.log
PGA_WS=13 timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2c_pytest_ws_all.log 2>&1; tail -4 gpurun_out/r2c_pytest_ws_all.log
timeout 300 python bench.py --steps 3 --warmup 3 --gens 100 > gpurun_out/r2c_bench_default.json 2> gpurun_out/r2c_bench_default.err; tail -c 1500 gpurun_out/r2c_bench_default.json
