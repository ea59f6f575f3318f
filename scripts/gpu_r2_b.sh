# round 2, GPU call B: extension tests, then the warp-specialised generation kernel:
# parity under PGA_WS, then timing of its four warp layouts
set -x
mkdir -p gpurun_out
timeout 900 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest.log 2>&1; tail -5 gpurun_out/r2b_pytest.log
PGA_WS=1 timeout 300 python -m pytest tests/test_gpu_parity.py -x -q -k "onemax" > gpurun_out/r2b_pytest_ws.log 2>&1; tail -5 gpurun_out/r2b_pytest_ws.log
show() { python -c "import json,sys; d=json.loads(sys.stdin.readline()); print('$1', 'us/gen', round(d['roofline']['avg_launch_ms']*1000,2),'frac', round(d['roofline']['frac'],3), 'evals/s %.4g'%d['value'])"; }
for ws in 0 1 2 3 4; do
  PGA_WS=$ws timeout 120 python bench.py --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax 2^20 ws=$ws"
done 2>&1 | tee gpurun_out/r2b_ws.log
for sz in 262144 4194304; do
  PGA_WS=1 timeout 120 python bench.py --size $sz --steps 5 --warmup 3 --gens 100 --no-cpu-baseline 2>/dev/null | show "onemax size=$sz ws=1"
done 2>&1 | tee -a gpurun_out/r2b_ws.log
PGA_WS=1 timeout 600 python -m pytest tests -m gpu -x -q > gpurun_out/r2b_pytest_ws_all.log 2>&1; tail -5 gpurun_out/r2b_pytest_ws_all.log
