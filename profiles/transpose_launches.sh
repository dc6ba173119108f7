#!/bin/bash
# per-launch times of the transpose kernels (20M-node workload); usage: bash profiles/transpose_launches.sh [out.csv]
out=${1:-gpurun_out/transpose_launches.csv}
ncu --metrics gpu__time_duration.sum,dram__bytes_read.sum,dram__bytes_write.sum,sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active,smsp__issue_active.avg.pct_of_peak_sustained_active,smsp__inst_executed.sum --clock-control none -k regex:"k_t_|k_rs_|k_scan|k_union" -c 24 --csv --log-file $out python bench.py --nodes 20000000 --steps 1 --warmup 1 --no-e2e --no-cpu-baseline --no-labels --no-configs > gpurun_out/tb_ncu.log 2>&1
python - $out <<'PY'
import csv,sys
rows=[r for r in csv.reader(open(sys.argv[1])) if len(r)>10]
h=rows[0]; ix={n:i for i,n in enumerate(h)}
cur={}
for r in rows[1:]:
    cur.setdefault((int(r[ix['ID']]), r[ix['Kernel Name']][:40]),{})[r[ix['Metric Name']]]=float(r[ix['Metric Value']].replace(',',''))
for k,v in sorted(cur.items()):
    if k[0] < 6: continue
    print(str(k[0]).rjust(3), k[1].ljust(40), "t=%.3fms R=%.2fGB W=%.2fGB adu=%.0f%% issue=%.0f%% inst=%.0fM" % (v['gpu__time_duration.sum']/1e6, v['dram__bytes_read.sum']/1e9, v['dram__bytes_write.sum']/1e9, v['sm__inst_executed_pipe_adu.avg.pct_of_peak_sustained_active'], v['smsp__issue_active.avg.pct_of_peak_sustained_active'], v['smsp__inst_executed.sum']/1e6))
PY
