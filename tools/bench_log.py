#!/usr/bin/env python
"""Print the bench JSON lines of a log (sections start with '== name') as one line each.  usage: bench_log.py file"""
import json
import sys
for l in open(sys.argv[1]):
    l = l.strip()
    if l.startswith('=='):
        print(l, end='  ')
    elif l.startswith('{'):
        d = json.loads(l)
        r = d.get("roofline") or {"frac": float("nan")}
        print(f"{d['value']/1e6:.4f} Mimg/s  {d['ms_per_step']*1e3:.1f} us  frac {r['frac']:.3f}  e2e {d['e2e']['value']/1e3:.0f}k")
    elif l:
        print(l[:200])
