#!/usr/bin/env python
"""Host <-> device copy bandwidth of this box with every GPU busy at once: the ceiling of bench.py's `e2e` leg.

    python tools/pcie_sweep.py                                         (1 GPU)
    python -m torch.distributed.run --nnodes=1 --nproc-per-node N --master-addr 127.0.0.1 --master-port 29511 tools/pcie_sweep.py

Every rank owns one GPU and moves pinned host memory: H2D alone, D2H alone, and both at once on two streams (the e2e
path uploads images while it downloads results).  Times are CUDA events per rank, the aggregate is the sum over ranks of
bytes / max-over-ranks time.  Prints one JSON line (rank 0) -> profiles/pcie_sweep.json collects them."""
import json
import os
import sys

import torch
import torch.distributed as dist


def main():
    rank, world, local = int(os.environ.get("RANK", 0)), int(os.environ.get("WORLD_SIZE", 1)), int(os.environ.get("LOCAL_RANK", 0))
    torch.cuda.set_device(local)
    dev = torch.device("cuda", local)
    if world > 1:
        dist.init_process_group("nccl", device_id=dev)
    nbytes = int(float(sys.argv[1]) * (1 << 20)) if len(sys.argv) > 1 else 256 << 20
    reps = 12
    h_up, h_dn = torch.empty(nbytes, dtype=torch.uint8).pin_memory(), torch.empty(nbytes, dtype=torch.uint8).pin_memory()
    d_up, d_dn = torch.empty(nbytes, dtype=torch.uint8, device=dev), torch.empty(nbytes, dtype=torch.uint8, device=dev)
    s_up, s_dn = torch.cuda.Stream(dev), torch.cuda.Stream(dev)

    def run(up, down):
        if world > 1:
            dist.barrier()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s_up.wait_stream(torch.cuda.current_stream()); s_dn.wait_stream(torch.cuda.current_stream())
        for _ in range(reps):
            if up:
                with torch.cuda.stream(s_up):
                    d_up.copy_(h_up, non_blocking=True)
            if down:
                with torch.cuda.stream(s_dn):
                    h_dn.copy_(d_dn, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s_up); torch.cuda.current_stream().wait_stream(s_dn)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1)
        if world > 1:
            t = torch.tensor([ms], device=dev)
            dist.all_reduce(t, op=dist.ReduceOp.MAX)
            ms = float(t.item())
        return ms

    run(True, True)
    res = {}
    for name, up, down in (("h2d", True, False), ("d2h", False, True), ("both", True, True)):
        ms = run(up, down)
        gb = nbytes * reps / 1e9
        res[name] = {"ms": ms, "per_gpu_GBps_each_direction": gb / (ms * 1e-3),
                     "aggregate_GBps": world * gb * (2 if (up and down) else 1) / (ms * 1e-3)}
    if rank == 0:
        print(json.dumps({"n_gpus": world, "bytes_per_copy": nbytes, "copies": reps, "host_cpus": len(os.sched_getaffinity(0)), **res}))
    if world > 1:
        dist.destroy_process_group()


if __name__ == "__main__":
    main()
