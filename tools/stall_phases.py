#!/usr/bin/env python
"""Stall reasons per phase of a v6 kernel capture (phases = SASS segments between the clock reads of SSAD_MARK), and
the instructions that collect the most samples in the two unit loops.   usage: stall_phases.py file.ncu-rep [n_warps]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
W = float(sys.argv[2]) if len(sys.argv) > 2 else 16384.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
cols = ['stall_barrier', 'stall_branch_resolving', 'stall_dispatch', 'stall_lg', 'stall_long_sb', 'stall_math', 'stall_membar', 'stall_mio',
        'stall_no_inst', 'stall_not_selected', 'stall_selected', 'stall_short_sb', 'stall_wait']
hd, data = None, []
for r in csv.reader(out.splitlines()):
    if r and r[0] == 'Address':
        hd = r; iS = hd.index('Source'); iI = hd.index('Instructions Executed'); iN = hd.index('# Samples'); ix = {c: hd.index(c) for c in cols}; continue
    if hd is None or len(r) <= iI:
        continue
    try:
        data.append((r[iS], int(r[iI]), int(r[iN] or 0), {c: int(r[ix[c]] or 0) for c in cols}))
    except ValueError:
        pass
marks = [i for i, d in enumerate(data) if 'SR_CLOCK' in d[0]]
names = ['prologue', 'geometry', 'pass1', 'exchange1', 'decide+canvas', 'mean+tables', 'pass2', 'tail']
bounds = [0] + marks + [len(data)]
print(f"{'phase':14s} {'instr/w':>8s} " + ' '.join(f'{c[6:][:8]:>8s}' for c in cols))
tot = collections.Counter()
for i in range(len(bounds) - 1):
    agg = collections.Counter()
    n = 0
    for d in data[bounds[i]:bounds[i + 1]]:
        n += d[1]
        for c in cols:
            agg[c] += d[3][c]
    tot.update(agg)
    print(f"{names[i] if i < len(names) else str(i):14s} {n / W:8.1f} " + ' '.join(f'{agg[c]:8d}' for c in cols))
print(f"{'total':14s} {'':8s} " + ' '.join(f'{tot[c]:8d}' for c in cols))
for i in (2, 6):
    if i + 1 >= len(bounds):
        continue
    seg = [(j,) + data[j] for j in range(bounds[i], bounds[i + 1])]
    seg.sort(key=lambda t: -t[3])
    print(names[i], 'instructions with most samples:')
    for t in seg[:16]:
        top = max(t[4].items(), key=lambda kv: kv[1])
        print(f'   {t[0]:6d} smp {t[3]:4d} ({top[0][6:]} {top[1]}) cnt/warp {t[2] / W:5.2f}  {t[1][:80]}')
