"""Summarises an ncu per-launch CSV collected with the metric list used in profiles/ (one line per launch)."""
import collections
import csv
import sys

lines = [l for l in open(sys.argv[1]) if not l.startswith('==')]
tab = collections.OrderedDict()
for row in csv.DictReader(lines):
    tab.setdefault((row['ID'], row['Kernel Name'][:40]), {})[row['Metric Name']] = row['Metric Value']


def f(m, k):
    return float(m.get(k, '0').replace(',', ''))


for (i, k), m in tab.items():
    print(f"{i:>4s} {k:40s} t={f(m, 'gpu__time_duration.sum') / 1e6:7.3f}ms inst={f(m, 'smsp__inst_executed.sum') / 1e6:8.1f}M "
          f"lanes={f(m, 'smsp__thread_inst_executed_per_inst_executed.ratio'):5.2f} warps={f(m, 'sm__warps_active.avg.pct_of_peak_sustained_active'):5.1f}% "
          f"issue={f(m, 'smsp__issue_active.avg.pct_of_peak_sustained_active'):5.1f}% dramR={f(m, 'dram__bytes_read.sum') / 1e9:5.2f}GB "
          f"W={f(m, 'dram__bytes_write.sum') / 1e9:5.2f}GB lts={f(m, 'lts__throughput.avg.pct_of_peak_sustained_elapsed'):4.1f}% "
          f"l1={f(m, 'l1tex__throughput.avg.pct_of_peak_sustained_active'):4.1f}%")
