# round-1 profiling pass (1 GPU): launch list of the default bench command, then one
# full capture of the fused kernel for OneMax and one for Rastrigin
set -x
CMD="python bench.py --steps 2 --warmup 3 --gens 10 --no-cpu-baseline"
$CMD > gpurun_out/plain_list.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/launches_r1.csv $CMD > gpurun_out/ncu_list.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --gens 5 --no-cpu-baseline"
$CMD > gpurun_out/plain_full.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:generation_fused_tiled -s 12 -c 2 -o gpurun_out/prof_onemax_r1 $CMD > gpurun_out/ncu_full_onemax.log 2>&1
CMD="python bench.py --workload rastrigin --steps 1 --warmup 3 --gens 3 --no-cpu-baseline"
$CMD > gpurun_out/plain_full2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:generation_fused_tiled -s 6 -c 2 -o gpurun_out/prof_rastrigin_r1 $CMD > gpurun_out/ncu_full_rastrigin.log 2>&1
CMD="python bench.py --workload test1 --steps 1 --warmup 3 --gens 5 --no-cpu-baseline"
$CMD > gpurun_out/plain_full3.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:generation_hooks -s 12 -c 2 -o gpurun_out/prof_test1_r1 $CMD > gpurun_out/ncu_full_test1.log 2>&1
tail -2 gpurun_out/ncu_full_*.log
