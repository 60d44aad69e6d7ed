#!/usr/bin/env python
"""Per-phase warp-instruction counts of a v6 kernel capture: segments the SASS of an .ncu-rep at the clock reads
(SSAD_MARK) and prints instructions per warp and the opcode mix of the two unit loops.
usage: phase_counts.py file.ncu-rep n_warps [units_per_warp]"""
import collections
import csv
import subprocess
import sys

rep, W = sys.argv[1], float(sys.argv[2])
upw = float(sys.argv[3]) if len(sys.argv) > 3 else 8.0
out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "sass", "--csv"], capture_output=True, text=True).stdout
hd, data = None, []
for r in csv.reader(out.splitlines()):
    if r and r[0] == 'Address':
        hd = r; iS = hd.index('Source'); iI = hd.index('Instructions Executed'); iN = hd.index('# Samples'); continue
    if hd is None or len(r) <= iI:
        continue
    try:
        data.append((r[iS], int(r[iI]), int(r[iN])))
    except ValueError:
        pass
tot = sum(d[1] for d in data)
print(len(data), 'sass instrs; executed per warp', round(tot / W, 1))
names = ['prologue', 'geometry', 'pass1', 'exchange1', 'decide+canvas', 'mean+tables', 'pass2', 'tail']
segs, cur = [], []
for d in data:
    if 'SR_CLOCK' in d[0]:
        segs.append(cur); cur = []
    cur.append(d)
segs.append(cur)
for i, sg in enumerate(segs):
    n = sum(d[1] for d in sg); s = sum(d[2] for d in sg)
    print(f'{names[i] if i < len(names) else i:14s} {n / W:8.1f} instr/warp   {100 * n / tot:5.1f}%   stall samples {s}')
for i in (2, 6):
    if i >= len(segs):
        continue
    ops = collections.Counter()
    for d in segs[i]:
        t = d[0].split(); ops[t[1] if t[0].startswith('@') else t[0]] += d[1]
    print(names[i], 'per unit:', round(sum(ops.values()) / W / upw, 1))
    print('   ' + '  '.join(f'{op} {n / W / upw:.1f}' for op, n in ops.most_common(24)))
