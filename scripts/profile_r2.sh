# round-2 profiling pass (1 GPU): launch lists of the default bench command and of an island run,
# full captures of the shipped fused kernel (OneMax, Rastrigin) and of the cooperative select
set -x
mkdir -p gpurun_out
CMD="python bench.py --steps 2 --warmup 3 --gens 10 --no-cpu-baseline --no-c5"
$CMD > gpurun_out/r2p_plain_list.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 400 --csv --log-file gpurun_out/r2_launches_onemax.csv $CMD > gpurun_out/r2p_ncu_list.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --gens 20 --no-cpu-baseline --no-c5 --islands 2"
$CMD > gpurun_out/r2p_plain_list2.log 2>&1 &&
ncu --metrics gpu__time_duration.sum --clock-control none -c 600 --csv --log-file gpurun_out/r2_launches_islands.csv $CMD > gpurun_out/r2p_ncu_list2.log 2>&1
CMD="python bench.py --steps 1 --warmup 3 --gens 5 --no-cpu-baseline --no-c5"
$CMD > gpurun_out/r2p_plain_full.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:generation_fused_tiled -s 12 -c 2 -o gpurun_out/prof_onemax_r2 $CMD > gpurun_out/r2p_ncu_full_onemax.log 2>&1
CMD="python bench.py --workload rastrigin --steps 1 --warmup 3 --gens 3 --no-cpu-baseline"
$CMD > gpurun_out/r2p_plain_full2.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:generation_fused_tiled -s 6 -c 2 -o gpurun_out/prof_rastrigin_r2 $CMD > gpurun_out/r2p_ncu_full_rastrigin.log 2>&1
CMD="python scripts/time_select.py 1048576"
$CMD > gpurun_out/r2p_plain_sel.log 2>&1 &&
ncu --set full --clock-control none --import-source on -k regex:topk_select_coop -s 4 -c 2 -o gpurun_out/prof_select_r2 $CMD > gpurun_out/r2p_ncu_full_select.log 2>&1
# gpurun brings back at most 64 MiB: summarise here, keep only the OneMax report
python scripts/ncu_summary.py gpurun_out/prof_onemax_r2.ncu-rep > gpurun_out/r2_ncu_onemax.txt 2>/dev/null
python scripts/ncu_summary.py gpurun_out/prof_rastrigin_r2.ncu-rep > gpurun_out/r2_ncu_rastrigin.txt 2>/dev/null
python scripts/ncu_summary.py gpurun_out/prof_select_r2.ncu-rep > gpurun_out/r2_ncu_select.txt 2>/dev/null
rm -f gpurun_out/prof_rastrigin_r2.ncu-rep gpurun_out/prof_select_r2.ncu-rep
python bench.py --steps 20 --warmup 3 > gpurun_out/r2_bench_default.json 2> gpurun_out/r2_bench_default.err; cut -c1-400 gpurun_out/r2_bench_default.json
python bench.py --impl reference --steps 3 --warmup 1 > gpurun_out/r2_bench_reference.json 2>/dev/null; cut -c1-300 gpurun_out/r2_bench_reference.json
for wl in rastrigin test1 tsp; do python bench.py --workload $wl --steps 3 --warmup 3 --gens 20 --no-cpu-baseline > gpurun_out/r2_bench_$wl.json 2>/dev/null; cut -c1-200 gpurun_out/r2_bench_$wl.json; done
python -c "import __graft_entry__ as g; g.smoke()"
