# per-kernel time and DRAM bytes of the first-parent-ordered Rastrigin generation against the child-ordered one
mkdir -p gpurun_out
for srt in 0 1; do
CMD="python bench.py --workload rastrigin --steps 1 --warmup 3 --gens 3 --no-cpu-baseline"
PGA_SORTED=$srt $CMD > gpurun_out/r2g_plain_$srt.log 2>&1 &&
PGA_SORTED=$srt ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,lts__t_sector_hit_rate.pct,smsp__issue_active.avg.pct_of_peak_sustained_active --clock-control none -s 30 -c 40 --csv --log-file gpurun_out/r2g_launches_sorted$srt.csv $CMD > gpurun_out/r2g_ncu_$srt.log 2>&1
done
python - <<'PY'
import csv
for srt in (0,1):
    rows=list(csv.reader(l for l in open(f'gpurun_out/r2g_launches_sorted{srt}.csv') if l.startswith('"')))
    hdr=rows[0]; ki=hdr.index('Kernel Name'); mi=hdr.index('Metric Name'); vi=hdr.index('Metric Value'); ii=hdr.index('ID')
    per={}
    for r in rows[1:]:
        per.setdefault((r[ii], r[ki][:60]), {})[r[mi]]=r[vi]
    for (i,k),m in list(per.items())[-14:]:
        print(srt, i, k, {a.split('__')[1][:18]:b for a,b in m.items()})
PY
