#!/usr/bin/env python
"""Aggregate an ncu --csv launch list (gpu__time_duration.sum [, smsp__inst_executed.sum]) per kernel.  usage: launch_list.py file.csv"""
import collections
import csv
import sys
rows = list(csv.reader(l for l in open(sys.argv[1]) if l.startswith('"')))
hd = rows[0]
ik, im, iv = hd.index('Kernel Name'), hd.index('Metric Name'), hd.index('Metric Value')
agg = collections.defaultdict(lambda: collections.defaultdict(list))
for r in rows[1:]:
    agg[r[ik].split('(')[0]][r[im]].append(float(r[iv].replace(',', '')))
tot = 0.0
for k, m in agg.items():
    t = m.get('gpu__time_duration.sum', [0])
    n = m.get('smsp__inst_executed.sum', [0])
    tot += sum(t) / len(t)
    print(f"{k:28s} launches {len(t):3d}  mean {sum(t)/len(t)/1e3:7.1f} us  min {min(t)/1e3:7.1f}  warp-instr {sum(n)/max(len(n),1)/1e6:7.2f} M")
print(f"sum of means {tot/1e3:.1f} us")
