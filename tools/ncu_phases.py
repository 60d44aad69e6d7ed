#!/usr/bin/env python
"""Per-phase shares of a kernel from an ncu report taken with --import-source on:
python tools/ncu_phases.py report.ncu-rep  -> instructions, stall samples and shared-memory wavefronts of every stretch of
SASS between two BAR.SYNC, with its top opcodes, plus the kernel's issue / data-pipe utilisation."""
import csv
import subprocess
import sys


def main():
    rep = sys.argv[1]
    out = subprocess.run(["ncu", "-i", rep, "--page", "source", "--csv", "--print-source", "sass"], capture_output=True, text=True).stdout
    rows = list(csv.reader(out.splitlines()))
    h = [i for i, r in enumerate(rows) if r and r[0] == "Address"][0]
    hdr = rows[h]
    ie, isamp, isrc = hdr.index("Instructions Executed"), hdr.index("# Samples"), hdr.index("Source")
    iw, ix = hdr.index("L1 Wavefronts Shared"), hdr.index("L1 Wavefronts Shared Excessive")
    segs, cur = [], {"inst": 0, "samp": 0, "wave": 0, "excess": 0, "ops": {}}
    for r in rows[h + 1:]:
        try:
            ex, sm = int(r[ie]), int(r[isamp])
        except ValueError:
            continue
        op = [t for t in r[isrc].split() if not t.startswith("@")][0]
        try:
            w, wx = int(r[iw]), int(r[ix])
        except ValueError:
            w = wx = 0
        cur["inst"] += ex; cur["samp"] += sm; cur["wave"] += w; cur["excess"] += wx
        cur["ops"][op.split(".")[0]] = cur["ops"].get(op.split(".")[0], 0) + ex
        if "BAR.SYNC" in r[isrc]:
            segs.append(cur)
            cur = {"inst": 0, "samp": 0, "wave": 0, "excess": 0, "ops": {}}
    segs.append(cur)
    ti, ts, tw = (sum(s[k] for s in segs) or 1 for k in ("inst", "samp", "wave"))
    print(f"warp instructions {ti}, shared-memory wavefronts {tw}")
    for i, s in enumerate(segs):
        top = sorted(s["ops"].items(), key=lambda kv: -kv[1])[:8]
        print(f"phase {i}: inst {100 * s['inst'] / ti:5.1f} %  samples {100 * s['samp'] / ts:5.1f} %  smem wavefronts {100 * s['wave'] / tw:5.1f} % "
              f"(excess {100 * s['excess'] / tw:4.1f})  " + " ".join(f"{k}:{100 * v / ti:.1f}" for k, v in top))
    raw = subprocess.run(["ncu", "-i", rep, "--page", "raw", "--csv"], capture_output=True, text=True).stdout
    rr = list(csv.reader(raw.splitlines()))
    d = dict(zip(rr[0], rr[-1]))
    for k in ("gpu__time_duration.sum", "sm__issue_active.avg.pct_of_peak_sustained_elapsed", "l1tex__data_pipe_lsu_wavefronts.avg.pct_of_peak_sustained_elapsed",
              "l1tex__data_pipe_lsu_wavefronts_mem_shared.sum.pct_of_peak_sustained_elapsed", "sm__pipe_fma_cycles_active.avg.pct_of_peak_sustained_elapsed",
              "sm__inst_executed_pipe_alu.avg.pct_of_peak_sustained_active", "sm__inst_executed_pipe_lsu.avg.pct_of_peak_sustained_active",
              "sm__warps_active.avg.pct_of_peak_sustained_active", "launch__registers_per_thread", "dram__bytes_read.sum", "dram__bytes_write.sum"):
        if k in d:
            print(f"{k} = {d[k]}")


if __name__ == "__main__":
    main()
