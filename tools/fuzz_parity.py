#!/usr/bin/env python
"""Randomised parity sweep of ssad_generate against oracle/generator.py (test infrastructure, like tests/):

    python tools/fuzz_parity.py [SECONDS] [SEED]        (on the GPU box)

Draws random configurations -- class policy (fixed-pose / texture / per-image mask), image sizes (width a multiple of 8,
any height), patch mode with random window sizes, area ratios, 3 or 4 labels, dense / sparse / empty masks -- and random
batches of records, runs the CUDA path and compares every image byte for byte (uint8 result, label, float32 normalised
result) with the CPU restatement.  Prints one line per configuration and a summary; rc 1 on any mismatch."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))
import numpy as np  # noqa: E402

import scenarios as S  # noqa: E402
from ssad_b200 import GeneratorConfig, sample_params, synth  # noqa: E402


def random_mask(rng, H, W):
    kind = rng.integers(0, 6)
    yy, xx = np.mgrid[0:H, 0:W]
    if kind == 0:
        return synth.disc_mask(H, W, float(rng.uniform(0.15, 0.6)))
    if kind == 1:                                   # off-centre ellipse (touches the border sometimes)
        cy, cx = rng.uniform(0, H), rng.uniform(0, W)
        return ((yy - cy) / (H * rng.uniform(0.1, 0.5))) ** 2 + ((xx - cx) / (W * rng.uniform(0.1, 0.5))) ** 2 < 1
    if kind == 2:                                   # sparse: a few hundred pixels
        return rng.random((H, W)) < rng.uniform(0.001, 0.02)
    if kind == 3:
        return np.zeros((H, W), dtype=bool)         # empty -> fallback centres
    if kind == 4:                                   # thin band
        return np.abs(yy - H * rng.uniform(0.2, 0.8)) < rng.integers(1, 6)
    return np.ones((H, W), dtype=bool)


def random_scenario(rng, k):
    policy = rng.choice(["fixed", "texture", "non_fixed"])
    subject = {"fixed": "bottle", "texture": "carpet", "non_fixed": "screw"}[str(policy)]
    W = int(rng.integers(8, 49)) * 8
    H = int(rng.integers(64, 385))
    if rng.random() < 0.5:
        H = W                                       # square images, as the reference uses
    batch = int(rng.integers(3, 25))
    patch = bool(rng.random() < 0.45)
    kw = dict(subject=subject, imsize=(W, H), num_labels=int(rng.choice([3, 4])))
    if patch:
        if subject == "screw":
            kw["subject"] = "hazelnut"              # (screw has a class pre-crop in patch mode: tests/scenarios.py screw_patch)
        kw["patch_localization"] = True
        kw["patch_size"] = int(rng.integers(4, min(W, H) // 8 + 1)) * 8
        if rng.random() < 0.5:
            kw["patch_area_ratio"] = (0.1, 0.2)
            kw["scar_area_ratio"] = (0.01, 0.02)
    elif rng.random() < 0.3:
        lo = float(rng.uniform(0.01, 0.2))
        kw["patch_area_ratio"] = (lo, lo + float(rng.uniform(0.0, 0.3)))
    donors = None
    if policy == "texture":
        nd = int(rng.integers(1, 16))
        kw["n_donors"] = nd
        donors = np.concatenate([synth.texture_images(1, H, W, 900 + 31 * k + j, tint=tuple(int(v) for v in rng.integers(30, 230, 3)))
                                 for j in range(nd)])
        imgs = synth.texture_images(batch, H, W, 7000 + k, tint=(120, 110, 100))
        mask = None
    elif policy == "fixed":
        imgs = synth.object_images(batch, H, W, 8000 + k)
        mask = random_mask(rng, H, W)
    else:
        imgs = synth.screw_images(batch, H, W, 9000 + k)
        mask = np.stack([random_mask(rng, H, W) for _ in range(batch)])
    cfg = GeneratorConfig(**kw)
    labels = rng.integers(0, cfg.num_labels, batch)
    params = sample_params(cfg, batch, np.random.default_rng(int(rng.integers(0, 1 << 31))), labels=labels)
    sc = S.Scenario(f"fuzz{k}", cfg, imgs, mask, donors, params)
    if np.any(labels == 3):
        params.resolve_lines(lambda i, sc=sc: S.window_mask(sc, i))
    return sc, kw


def main():
    import gpu_utils as U
    from oracle import generator as og
    from PIL import Image
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rng = np.random.default_rng(seed)
    t0 = time.time()
    k = nimg = bad = 0
    while time.time() - t0 < budget:
        sc, kw = random_scenario(rng, k)
        k += 1
        try:
            out = U.run_cuda(sc)
        except Exception as e:                      # a configuration the library refuses: report, it is not a mismatch
            print(f"[{k}] {kw} B={len(sc.params)}: refused: {type(e).__name__}: {e}")
            continue
        x, y, xn = out["x_u8"].cpu().numpy(), out["y"].cpu().numpy(), out["x"].cpu().numpy()
        nb = 0
        for i in range(len(sc.params)):
            ox, oy, _ = S.run_oracle(sc, i)
            ok = int(y[i]) == oy and np.array_equal(x[i], ox)
            if ok and i < 3:
                ok = np.array_equal(xn[i], og.to_tensor_normalize(Image.fromarray(ox)).numpy())
            if not ok:
                nb += 1
                d = np.abs(x[i].astype(np.int16) - ox.astype(np.int16))
                print(f"   MISMATCH image {i}: label {int(y[i])} vs {oy}, {int((d > 0).sum())} bytes differ (max {int(d.max())})")
        nimg += len(sc.params)
        bad += nb
        print(f"[{k}] {kw} B={len(sc.params)} labels={np.bincount(sc.params.draws['label'], minlength=4).tolist()}: "
              f"{'ok' if nb == 0 else f'{nb} MISMATCHES'}", flush=True)
    print(f"fuzz: {k} configurations, {nimg} images, {bad} mismatches, seed {seed}, {time.time() - t0:.0f} s")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
