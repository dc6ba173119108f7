#!/bin/bash
# per-launch times and DRAM bytes of one decode step; usage: bash profiles/step_launches.sh NODES out.csv
nodes=${1:-20000000}; out=${2:-gpurun_out/step_launches.csv}
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,smsp__inst_executed.sum,smsp__thread_inst_executed_per_inst_executed.ratio,smsp__issue_active.avg.pct_of_peak_sustained_active,sm__warps_active.avg.pct_of_peak_sustained_active --clock-control none --csv --log-file $out python bench.py --nodes $nodes --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-labels --no-configs --no-transpose > gpurun_out/step_ncu.log 2>&1
python - $out <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
h=rows[0]; ix={n:i for i,n in enumerate(h)}
cur={}
for r in rows[1:]:
    cur.setdefault((int(r[ix['ID']]), r[ix['Kernel Name']][:44]),{})[r[ix['Metric Name']]]=float(r[ix['Metric Value']].replace(',',''))
items=sorted(cur.items())
# the last step = the launches after the last k_extras-start
starts=[i for i,(k,v) in enumerate(items) if k[1].startswith('void k_extras<') or k[1].startswith('k_extras<')]
last=starts[-1] if starts else 0
tot=0; R=0; W=0
for k,v in items[last:]:
    t=v['gpu__time_duration.sum']/1e6; tot+=t; R+=v['dram__bytes_read.sum']; W+=v['dram__bytes_write.sum']
    print(str(k[0]).rjust(4), k[1].ljust(44), "t=%7.3fms inst=%8.1fM lanes=%5.2f warps=%5.1f%% issue=%5.1f%% dramR=%6.2fGB W=%6.2fGB" % (t, v['smsp__inst_executed.sum']/1e6, v['smsp__thread_inst_executed_per_inst_executed.ratio'], v['sm__warps_active.avg.pct_of_peak_sustained_active'], v['smsp__issue_active.avg.pct_of_peak_sustained_active'], v['dram__bytes_read.sum']/1e9, v['dram__bytes_write.sum']/1e9))
print("sum of launches %.3f ms; DRAM read %.2f GB + written %.2f GB = %.2f GB" % (tot, R/1e9, W/1e9, (R+W)/1e9))
PY
