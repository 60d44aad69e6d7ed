#!/usr/bin/env python
"""Warp instructions and stall samples per CUDA source line of one kernel of an ncu capture.
The SASS page of the report gives per-instruction counts; nvdisasm --print-line-info of the same build's cubin gives
the source line of every instruction (same order).   usage: line_profile.py sass.csv kernel_index cubin kernel_name [top [samples]]"""
import collections, csv, re, subprocess, sys
sass, ki, cubin, kname = sys.argv[1], int(sys.argv[2]), sys.argv[3], sys.argv[4]
top = int(sys.argv[5]) if len(sys.argv) > 5 else 40
rows = list(csv.reader(open(sass)))
starts = [i for i, r in enumerate(rows) if r and r[0] == 'Address'] + [len(rows)]
hd = rows[starts[ki]]; iI = hd.index('Instructions Executed'); iN = hd.index('# Samples'); iS = hd.index('Source')
data = [(r[iS], int(r[iI]), int(r[iN])) for r in rows[starts[ki] + 1:starts[ki + 1]] if len(r) > iI and r[iI].isdigit()]
dis = subprocess.run(['nvdisasm', '--print-line-info', cubin], capture_output=True, text=True).stdout.splitlines()
lines, cur, on = [], None, False
for l in dis:
    if l.startswith('.text.'):
        on = kname in l
        continue
    if not on:
        continue
    m = re.search(r'//## File "([^"]+)", line (\d+)', l)
    if m:
        cur = (m.group(1).split('/')[-1], int(m.group(2))); continue
    if re.match(r'\s+/\*[0-9a-f]{4,}\*/\s+\S', l):
        lines.append(cur)
if len(lines) != len(data):
    print('warning: instruction counts differ', len(lines), len(data))
agg = collections.defaultdict(lambda: [0, 0, 0])
for (src, ex, n), ln in zip(data, lines):
    a = agg[ln]; a[0] += ex; a[1] += n; a[2] += 1
tot = sum(d[1] for d in data); totn = sum(d[2] for d in data)
print('total warp-instr', tot, 'samples', totn)
srcs = {}
bysmp = len(sys.argv) > 6 and sys.argv[6] == "samples"
for (f, ln), (ex, n, k) in sorted(agg.items(), key=lambda x: -x[1][1 if bysmp else 0])[:top]:
    if f not in srcs:
        try: srcs[f] = open('self-supervised-anomaly-detection_b200/csrc/' + f).read().splitlines()
        except OSError: srcs[f] = []
    text = srcs[f][ln - 1].strip()[:90] if ln - 1 < len(srcs[f]) else ''
    print(f"{100.0*ex/tot:5.1f}% instr {100.0*n/max(totn,1):5.1f}% smp  {f.replace('ssad_','')[:14]:14s}:{ln:4d}  {text}")
