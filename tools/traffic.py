#!/usr/bin/env python
"""Whole-step DRAM traffic of a bench configuration, for `roofline.traffic`.

    python tools/traffic.py CONFIG [OUT] [STEPS]      (on the GPU box; writes / updates profiles/traffic.json)

Runs  ncu --profile-from-start off --cache-control none --clock-control none --metrics dram__bytes_read.sum,
dram__bytes_write.sum,gpu__time_duration.sum  over  bench.py --config CONFIG --profile-steps STEPS --graph 0: every kernel
of STEPS consecutive steps, in issue order, WITHOUT a cache flush between kernels -- so a re-read that hits the L2 is not
counted as DRAM traffic, and lines written by one kernel and evicted during a later one are counted where they leave
the L2.  Reads and writes of all kernels are summed and divided by STEPS (several steps: the write-backs of a step that
are still in the L2 when it ends are caught by the next one)."""
import csv
import json
import os
import subprocess
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def main():
    config = int(sys.argv[1])
    out = sys.argv[2] if len(sys.argv) > 2 else "u8"
    steps = int(sys.argv[3]) if len(sys.argv) > 3 else 4
    log = os.path.join(ROOT, "gpurun_out", f"traffic_config{config}_{out}.csv")
    os.makedirs(os.path.dirname(log), exist_ok=True)
    cmd = ["ncu", "--profile-from-start", "off", "--cache-control", "none", "--clock-control", "none", "--metrics",
           "dram__bytes_read.sum,dram__bytes_write.sum,gpu__time_duration.sum", "--csv", "--log-file", log,
           sys.executable, os.path.join(ROOT, "bench.py"), "--config", str(config), "--out", out, "--graph", "0", "--inflight", "1",
           "--profile-steps", str(steps), "--no-cpu", "--per-config", "0", "--verify", "0"]
    subprocess.run(cmd, check=True, stdout=subprocess.DEVNULL)
    rows = list(csv.reader(l for l in open(log) if l.startswith('"')))
    head = rows[0]
    ik, im, iv, iu = head.index("Kernel Name"), head.index("Metric Name"), head.index("Metric Value"), head.index("Metric Unit")
    scale = {"byte": 1, "Kbyte": 1e3, "Mbyte": 1e6, "Gbyte": 1e9, "ns": 1e-3, "us": 1.0, "ms": 1e3, "usecond": 1.0, "nsecond": 1e-3, "msecond": 1e3}
    tot = {"dram__bytes_read.sum": 0.0, "dram__bytes_write.sum": 0.0, "gpu__time_duration.sum": 0.0}
    per_kernel = {}
    for r in rows[1:]:
        v = float(r[iv].replace(",", "")) * scale.get(r[iu], 1)
        tot[r[im]] += v
        name = r[ik].split("(")[0].split("::")[-1]
        per_kernel.setdefault(name, {}).setdefault(r[im], 0.0)
        per_kernel[name][r[im]] += v
    res = {"read_bytes_per_step": tot["dram__bytes_read.sum"] / steps, "write_bytes_per_step": tot["dram__bytes_write.sum"] / steps,
           "bytes_per_step": (tot["dram__bytes_read.sum"] + tot["dram__bytes_write.sum"]) / steps,
           "kernel_us_per_step_serialised": tot["gpu__time_duration.sum"] / steps, "steps": steps,
           "per_kernel_mb_per_step": {k: {"read": v.get("dram__bytes_read.sum", 0) / steps / 1e6, "write": v.get("dram__bytes_write.sum", 0) / steps / 1e6,
                                          "us": v.get("gpu__time_duration.sum", 0) / steps} for k, v in per_kernel.items()},
           "how": "ncu --cache-control none, all kernels of consecutive steps (plain launches, one stream), summed / steps"}
    path = os.path.join(ROOT, "profiles", "traffic.json")
    try:
        allt = json.load(open(path))
    except Exception:
        allt = {}
    allt[f"config{config}_{out}"] = res["bytes_per_step"]
    allt.setdefault("detail", {})[f"config{config}_{out}"] = res
    json.dump(allt, open(path, "w"), indent=1)
    json.dump(allt, open(os.path.join(ROOT, "gpurun_out", "traffic.json"), "w"), indent=1)      # (gpurun merges gpurun_out/ back)
    print(json.dumps({k: v for k, v in res.items() if k != "per_kernel_mb_per_step"}))
    for k, v in res["per_kernel_mb_per_step"].items():
        print(f"  {k:28s} read {v['read']:9.2f} MB  write {v['write']:9.2f} MB  {v['us']:9.1f} us")


if __name__ == "__main__":
    main()
