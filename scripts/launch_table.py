"""Aggregate an ncu launch list (--csv, gpu__time_duration.sum [+ smsp__inst_executed.sum]) per kernel.  Usage: launch_table.py CSV"""
import collections
import csv
import sys
rows = list(csv.reader(open(sys.argv[1])))
for i, r in enumerate(rows):
    if 'Kernel Name' in r:
        hdr, start = r, i + 1
        break
kn, mn, mv = hdr.index('Kernel Name'), hdr.index('Metric Name'), hdr.index('Metric Value')
agg = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[start:]:
    if len(r) > mv:
        agg[r[kn][:80]][r[mn]].append(float(r[mv].replace(',', '')))
for k, d in agg.items():
    t = d.get('gpu__time_duration.sum', [0])
    i = d.get('smsp__inst_executed.sum', [0])
    print(f"{k:80s} n={len(t):3d} mean={sum(t)/len(t)/1000:9.1f} us  inst={sum(i)/max(len(i),1)/1e6:8.2f} M")
