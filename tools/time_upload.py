#!/usr/bin/env python
"""Times ssad_upload_windows (kernel reads of pinned host memory) against a plain pinned copy of the whole batch:
python tools/time_upload.py [batch] [S] [ps]"""
import ctypes as C
import os
import sys

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import torch

from ssad_b200 import GeneratorConfig, _native, sample_params


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 512
    S = int(sys.argv[2]) if len(sys.argv) > 2 else 320
    ps = int(sys.argv[3]) if len(sys.argv) > 3 else 256
    lib = _native.load()
    cfg = GeneratorConfig(subject="bottle", imsize=(S, S), patch_localization=True, patch_size=ps, patch_area_ratio=(0.1, 0.2),
                          scar_area_ratio=(0.01, 0.02))
    p = sample_params(cfg, B, np.random.default_rng(0))
    rec = np.ascontiguousarray(p.records)
    d_rec = torch.from_numpy(rec.view(np.uint8).reshape(B, -1)).cuda()
    h = torch.randint(0, 255, (B, S, S, 3), dtype=torch.uint8).pin_memory()
    d = torch.zeros((B, S, S, 3), dtype=torch.uint8, device="cuda")
    st = torch.cuda.current_stream().cuda_stream
    max_h = int(rec["rect"][:, 3].max())

    def win():
        rc = lib.ssad_upload_windows(B, C.c_void_p(d_rec.data_ptr()), C.c_void_p(h.data_ptr()), C.c_void_p(d.data_ptr()), S * 3, S * S * 3,
                                     S, S, ps, ps, max_h, C.c_void_p(st))
        assert rc == 0, rc

    def full():
        d.copy_(h, non_blocking=True)
    for fn, name, nbytes in ((full, "full copy", B * S * S * 3), (win, "windows", B * ps * ps * 3)):
        for _ in range(3):
            fn()
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        for _ in range(20):
            fn()
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"{name}: {ms:.3f} ms, {nbytes / ms / 1e6:.1f} GB/s of useful bytes ({nbytes / 1e6:.1f} MB)")
    # the same two uploads with a device-to-host copy of the results' size running on another stream
    out_d = torch.zeros((B, ps, ps, 3), dtype=torch.uint8, device="cuda")
    out_h = torch.empty((B, ps, ps, 3), dtype=torch.uint8).pin_memory()
    s2 = torch.cuda.Stream()
    for fn, name, nbytes in ((full, "full copy", B * S * S * 3), (win, "windows", B * ps * ps * 3)):
        torch.cuda.synchronize()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record()
        s2.wait_stream(torch.cuda.current_stream())
        for _ in range(20):
            fn()
            with torch.cuda.stream(s2):
                out_h.copy_(out_d, non_blocking=True)
        torch.cuda.current_stream().wait_stream(s2)
        e1.record()
        torch.cuda.synchronize()
        ms = e0.elapsed_time(e1) / 20
        print(f"{name} + concurrent D2H of {out_d.numel() / 1e6:.1f} MB: {ms:.3f} ms per pair")
    # the windows really arrived
    i = B // 2
    x, y = int(rec["crop_off"][i, 0]), int(rec["crop_off"][i, 1])
    assert torch.equal(d[i, y:y + ps, x:x + ps].cpu(), h[i, y:y + ps, x:x + ps])


if __name__ == "__main__":
    main()
