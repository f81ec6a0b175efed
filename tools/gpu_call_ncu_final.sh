#!/bin/bash
# refresh of the headline kernel's ncu summary and of the launch list of the default bench command, on the final build
mkdir -p gpurun_out
timeout 600 ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 -o /tmp/atrium_final python bench.py --steps 2 --warmup 3 --no-cpu-baseline > gpurun_out/ncu_final.log 2>&1
python tools/ncu_summary.py /tmp/atrium_final.ncu-rep gpurun_out/r02_ncu_atrium_final.json "ncu --set full --clock-control none --import-source on -k regex:render_kernel -s 4 -c 1 python bench.py --steps 2 --warmup 3 --no-cpu-baseline" > /dev/null 2>&1
rm -f /tmp/atrium_final.ncu-rep
timeout 600 ncu --metrics gpu__time_duration.sum --clock-control none -c 4000 --csv --log-file gpurun_out/r02_launches_atrium_final.csv python bench.py --steps 20 --warmup 3 --no-cpu-baseline > gpurun_out/launches_final.log 2>&1
python - <<'PY'
import json, csv, collections
d=json.load(open('gpurun_out/r02_ncu_atrium_final.json'))['launches'][0]
print({k:d.get(k) for k in ('gpu__time_duration.sum','launch__registers_per_thread','smsp__issue_active.avg.pct_of_peak_sustained_active','smsp__thread_inst_executed_per_inst_executed.ratio','l1tex__t_sector_hit_rate.pct','lts__t_sector_hit_rate.pct','dram__bytes_read.sum','dram__bytes_write.sum','sm__warps_active.avg.pct_of_peak_sustained_active','smsp__inst_executed.sum')})
rows=[r for r in csv.reader(open('gpurun_out/r02_launches_atrium_final.csv')) if len(r)>5]
for i,r in enumerate(rows):
    if r[0]=='ID': hdr=r; rows=rows[i+1:]; break
ki=hdr.index('Kernel Name'); vi=hdr.index('Metric Value')
agg=collections.defaultdict(lambda:[0,0.0])
for r in rows:
    k=r[ki].split('(')[0][-40:]; v=float(r[vi].replace(',',''))/1000
    agg[k][0]+=1; agg[k][1]+=v
tot=sum(a[1] for a in agg.values())
w=csv.writer(open('gpurun_out/r02_launches_atrium_final_by_kernel.csv','w'))
w.writerow(['kernel','launches','total_us','share'])
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1]):
    w.writerow([k,n,round(t,1),round(t/tot,4)])
for k,(n,t) in sorted(agg.items(), key=lambda x:-x[1][1])[:8]: print(f"{k:42s} n={n:4d} total={t:9.1f} us share={t/tot:.3f}")
PY
