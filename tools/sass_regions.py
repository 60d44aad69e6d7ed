#!/usr/bin/env python
"""Regions of a kernel's SASS by execution count: consecutive instructions with the same 'Instructions Executed' are
merged; prints start, length, executions, share of all warp instructions, samples.   usage: sass_regions.py sass.csv [kernel_index] [min_share]"""
import csv, sys
rows = list(csv.reader(open(sys.argv[1])))
ki = int(sys.argv[2]) if len(sys.argv) > 2 else 0
ms = float(sys.argv[3]) if len(sys.argv) > 3 else 0.5
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address']
s = starts[ki]; e = starts[ki + 1] if ki + 1 < len(starts) else len(rows)
hd = rows[s]; iS = hd.index('Source'); iI = hd.index('Instructions Executed'); iN = hd.index('# Samples')
data = []
for r in rows[s + 1:e]:
    try: data.append((r[iS], int(r[iI]), int(r[iN])))
    except (ValueError, IndexError): pass
tot = sum(d[1] for d in data); totn = sum(d[2] for d in data)
print(len(data), 'instrs', tot, 'warp-instr', totn, 'samples')
i = 0
while i < len(data):
    j = i
    while j + 1 < len(data) and abs(data[j + 1][1] - data[i][1]) <= 0.02 * max(data[i][1], 1): j += 1
    n = j - i + 1; ex = sum(d[1] for d in data[i:j + 1]); sm = sum(d[2] for d in data[i:j + 1])
    if 100.0 * ex / tot >= ms:
        print(f"{i:5d} +{n:4d}  exec {data[i][1]:9d}  share {100.0*ex/tot:5.1f}%  samples {100.0*sm/max(totn,1):5.1f}%   {data[i][0][:60]}")
    i = j + 1
