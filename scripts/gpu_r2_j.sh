mkdir -p gpurun_out
python scripts/dbg_select.py 20000; python scripts/dbg_select.py 1048576; PGA_NO_COOP=1 python scripts/dbg_select.py 1048576
python scripts/dbg_select.py 1048576 > /dev/null 2>&1 && ncu --set full --clock-control none --import-source on -k regex:topk_select_coop -s 4 -c 2 -o gpurun_out/prof_select_r2 python scripts/dbg_select.py 1048576 > gpurun_out/r2j_ncu.log 2>&1; tail -2 gpurun_out/r2j_ncu.log
