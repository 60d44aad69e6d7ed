#!/usr/bin/env python
"""Host-side profile (cProfile) of the user-facing calls at a small batch, where the host, not the GPU, sets the rate:
python tools/profile_host.py [batch]   ->  BatchLoader epoch (datamodule path) and generate_from_host (e2e path)."""
import cProfile
import os
import pstats
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from ssad_b200 import GeneratorConfig, sample_params, synth  # noqa: E402
from ssad_b200.datasets import BatchLoader, PretextTaskDataset  # noqa: E402
from ssad_b200.generator import BatchGenerator  # noqa: E402


def main():
    B = int(sys.argv[1]) if len(sys.argv) > 1 else 96
    S = 256
    base = synth.object_images(40, S, S, 4242)
    ds = PretextTaskDataset("bottle", imsize=(S, S), images=base, fixed_segmentation=synth.disc_mask(S, S, 0.4), device="cuda", seed=7,
                            dataset_dir="/nonexistent/")
    ds.images_filenames = np.arange(1000)
    ds.index = np.arange(1000) % 40
    loader = BatchLoader(ds, B, shuffle=True, drop_last=True)
    for _ in loader:
        pass
    torch.cuda.synchronize()
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable()
    n = 0
    for _ in range(5):
        for b in loader:
            n += int(b[1].shape[0])
    torch.cuda.synchronize()
    pr.disable()
    print(f"== BatchLoader: {n / (time.perf_counter() - t0):.0f} img/s under the profiler")
    pstats.Stats(pr).sort_stats("cumulative").print_stats(35)

    cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
    gen = BatchGenerator(cfg, "cuda", mask=synth.disc_mask(S, S, 0.4))
    imgs = base[np.arange(B) % 40]
    h_src = torch.from_numpy(imgs).pin_memory()
    h_out = torch.empty((B, S, S, 3), dtype=torch.uint8).pin_memory()
    h_y = torch.empty((B,), dtype=torch.int64).pin_memory()
    rng = np.random.default_rng(0)
    ps = [sample_params(cfg, B, rng) for _ in range(50)]
    for p in ps[:3]:
        gen.generate_from_host(h_src, p, h_out, h_y)
    pr = cProfile.Profile()
    t0 = time.perf_counter()
    pr.enable()
    for p in ps:
        gen.generate_from_host(h_src, p, h_out, h_y)
    pr.disable()
    print(f"== generate_from_host: {B * len(ps) / (time.perf_counter() - t0):.0f} img/s under the profiler (records drawn ahead)")
    pstats.Stats(pr).sort_stats("cumulative").print_stats(35)


if __name__ == "__main__":
    main()
