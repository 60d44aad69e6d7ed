#!/usr/bin/env python
"""Times ssad_obj_mask on the config-4 images (screw-like, 1024x1024): python tools/time_objmask.py [batch] [steps]."""
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

import ssad_b200
from ssad_b200 import synth
from ssad_b200.generator import obj_mask_batch


def main():
    batch = int(sys.argv[1]) if len(sys.argv) > 1 else 256
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    S = int(sys.argv[3]) if len(sys.argv) > 3 else 1024
    base = synth.screw_images(8, S, S, 1238)
    src = torch.from_numpy(base[np.arange(batch) % 8]).cuda()
    lib = ssad_b200._native.load()
    ws = torch.empty((int(lib.ssad_obj_mask_workspace(batch, S, S)) + 16,), dtype=torch.uint8, device="cuda")
    out = obj_mask_batch(src, workspace=ws)
    torch.cuda.synchronize()
    for _ in range(2):
        obj_mask_batch(src, out=out, workspace=ws)
    torch.cuda.synchronize()
    e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    e0.record()
    for _ in range(steps):
        obj_mask_batch(src, out=out, workspace=ws)
    e1.record()
    torch.cuda.synchronize()
    ms = e0.elapsed_time(e1) / steps
    print(f"obj_mask: batch {batch} of {S}x{S}: {ms:.3f} ms/call = {ms * 1024 / batch:.3f} ms per 1024 images; "
          f"mask area mean {float(torch.tensor([bin(int(v) & 0xFFFFFFFF).count('1') for v in out[0].flatten().tolist()]).sum()):.0f} px")


if __name__ == "__main__":
    main()
