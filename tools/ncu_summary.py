#!/usr/bin/env python
"""Summarise an .ncu-rep (first profiled launch): key raw metrics, SASS opcode mix, and the CUDA source
lines ranked by instructions executed / stall samples.   usage: ncu_summary.py file.ncu-rep [topN [launch_index]]"""
import collections
import csv
import subprocess
import sys

rep = sys.argv[1]
topn = int(sys.argv[2]) if len(sys.argv) > 2 else 30
skip = ["--launch-skip", sys.argv[3], "--launch-count", "1"] if len(sys.argv) > 3 else []   # which profiled launch
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"] + skip, capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
h, units, v = r[0], r[1], r[2]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__cluster_size', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__cycles_elapsed.avg', 'smsp__cycles_active.avg', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum', 'lts__t_bytes.sum', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_fp64.avg.pct_of_peak_sustained_active', 'sm__pipe_fp64_cycles_active.avg.pct_of_peak_sustained_active',
        'launch__occupancy_limit_registers', 'launch__occupancy_limit_shared_mem', 'launch__occupancy_limit_warps',
        'launch__shared_mem_per_block_dynamic', 'dram__throughput.avg.pct_of_peak_sustained_elapsed']
print("kernel:", v[h.index('Kernel Name')] if 'Kernel Name' in h else '?')
for i, k in enumerate(h):
    if k in keys or ('issue_stalled' in k and 'per_issue_active' in k and 'not_issued' not in k):
        try:
            if 'issue_stalled' in k and float(v[i]) < 0.05:
                continue
        except ValueError:
            pass
        print(f"{k:95s} {units[i]:12s} {v[i]}")

src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--print-source", "cuda,sass", "--csv"] + skip,
                     capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
ops, smp = collections.Counter(), collections.Counter()
lines = {}          # (file, line) -> [text, instr, samples]
tot = tots = 0
cur_file, hd, nkern, seen_hdr = None, None, 0, 0
insts = []
for row in rows:
    if not row:
        continue
    if row[0] == "File Path":
        cur_file = row[1].split("/")[-1]
        continue
    if row[0] == "Function Name":
        continue
    if row[0] == "Kernel Name":
        nkern += 1
        if nkern > 1:
            break
        continue
    if row[0] == "Line No":
        hd = row
        iA = hd.index("Address")
        iS = iA + 1
        iI, iN = hd.index("Instructions Executed"), hd.index("# Samples")
        continue
    if hd is None or len(row) <= iI:
        continue
    if row[0] != "":            # a CUDA source line with aggregated metrics
        try:
            key = (cur_file, int(row[0]))
            ent = lines.setdefault(key, [row[1].strip(), 0, 0])
            ent[1] += int(row[iI]); ent[2] += int(row[iN])
        except ValueError:
            pass
        continue
    if row[iA] in ("", "...", "-"):
        continue
    t = row[iS].split()
    if not t:
        continue
    op = (t[1] if t[0].startswith('@') and len(t) > 1 else t[0]).split('.')[0]
    try:
        n, s = int(row[iI]), int(row[iN])
    except ValueError:
        continue
    ops[op] += n; smp[op] += s; tot += n; tots += s
    insts.append((s, n, row[iS].strip(), cur_file))
print('total warp instr (sass rows shown)', tot, 'samples', tots)
for op, n in ops.most_common(26):
    print(f'  {op:10s} {n:12d} {100*n/max(tot,1):5.1f}%  samples {smp[op]:7d} {100*smp[op]/max(tots,1):5.1f}%')
ltot = sum(e[1] for e in lines.values()); lsm = sum(e[2] for e in lines.values())
print(f'source lines by warp instructions (total {ltot}, samples {lsm}):')
for (f, ln), e in sorted(lines.items(), key=lambda kv: -kv[1][1])[:topn]:
    print(f'  {100*e[1]/max(ltot,1):5.1f}% instr {100*e[2]/max(lsm,1):5.1f}% smp  {f}:{ln:<4d} {e[0][:100]}')
print('source lines by stall samples:')
for (f, ln), e in sorted(lines.items(), key=lambda kv: -kv[1][2])[:topn]:
    print(f'  {100*e[2]/max(lsm,1):5.1f}% smp {100*e[1]/max(ltot,1):5.1f}% instr  {f}:{ln:<4d} {e[0][:100]}')
