#!/usr/bin/env python
"""Summarise an .ncu-rep: key raw metrics + SASS opcode mix + top stall lines.  usage: ncu_summary.py file.ncu-rep"""
import collections, csv, subprocess, sys
rep = sys.argv[1]
raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
r = list(csv.reader(raw.splitlines()))
h, units, v = r[0], r[1], r[2]
keys = ['gpu__time_duration.sum', 'dram__bytes_read.sum', 'dram__bytes_write.sum', 'gpu__dram_throughput.avg.pct_of_peak_sustained_elapsed',
        'sm__throughput.avg.pct_of_peak_sustained_elapsed', 'sm__warps_active.avg.pct_of_peak_sustained_active', 'launch__registers_per_thread',
        'launch__grid_size', 'launch__block_size', 'launch__cluster_size', 'smsp__inst_executed.sum', 'smsp__issue_active.avg.pct_of_peak_sustained_active',
        'sm__cycles_elapsed.avg', 'smsp__cycles_active.avg', 'l1tex__data_bank_conflicts_pipe_lsu_mem_shared.sum',
        'lts__t_sectors_op_read.sum', 'lts__t_sectors_op_write.sum', 'sm__inst_executed_pipe_fma.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_fmaheavy.avg.pct_of_peak_sustained_active',
        'sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_active', 'sm__pipe_alu_cycles_active.avg.pct_of_peak_sustained_active',
        'sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active', 'sm__inst_executed_pipe_xu.avg.pct_of_peak_sustained_active']
for i, k in enumerate(h):
    if k in keys or 'issue_stalled' in k and 'per_issue_active' in k and 'not_issued' not in k:
        try:
            if 'issue_stalled' in k and float(v[i]) < 0.05: continue
        except ValueError: pass
        print(f"{k:95s} {units[i]:12s} {v[i]}")
src = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv"], capture_output=True, text=True).stdout
rows = list(csv.reader(src.splitlines()))
hd = rows[1]; data = rows[2:]
iS, iI, iN = hd.index('Source'), hd.index('Instructions Executed'), hd.index('# Samples')
ops, smp = collections.Counter(), collections.Counter(); tot = 0; tots = 0
for row in data:
    if len(row) <= iI: continue
    t = row[iS].split()
    op = (t[1] if t[0].startswith('@') else t[0]).split('.')[0]
    n = int(row[iI]); ops[op] += n; smp[op] += int(row[iN]); tot += n; tots += int(row[iN])
print('total warp instr', tot, 'static', len(data), 'samples', tots)
for op, n in ops.most_common(24): print(f'  {op:10s} {n:12d} {100*n/tot:5.1f}%  samples {smp[op]:7d} {100*smp[op]/max(tots,1):5.1f}%')
print('top stall instructions:')
for row in sorted(data, key=lambda q: -int(q[iN]))[:14]:
    print(f'  {row[iN]:>7s} x{row[iI]:>9s}  {row[iS][:90]}')
