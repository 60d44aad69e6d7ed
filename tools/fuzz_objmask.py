#!/usr/bin/env python
"""Randomised parity sweep of ssad_obj_mask against oracle/objmask.py (test infrastructure, like tests/):

    python tools/fuzz_objmask.py [SECONDS] [SEED]        (on the GPU box)

Random image kinds (screw-like, object, texture, iid noise, blocks, smooth ramps, flat), random sizes (13 <= W <= 1024, any
H >= 13), random batches; the masks must be bit-identical.  rc 1 on any mismatch."""
import os
import sys
import time

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np  # noqa: E402
import torch  # noqa: E402

from ssad_b200 import synth  # noqa: E402
from ssad_b200.generator import obj_mask_batch, unpack_mask_bits  # noqa: E402


def random_images(rng, k):
    H = int(rng.integers(13, 400))
    W = int(rng.integers(13, 420))
    if rng.random() < 0.1:
        W = int(rng.integers(900, 1025))
        H = int(rng.integers(13, 200))
    n = int(rng.integers(1, 5))
    kind = int(rng.integers(0, 7))
    if kind == 0 and min(H, W) >= 32:
        img = synth.screw_images(n, H, W, 100 + k)
    elif kind == 1 and min(H, W) >= 32:
        img = synth.object_images(n, H, W, 200 + k)
    elif kind == 2 and min(H, W) >= 32:
        img = synth.texture_images(n, H, W, 300 + k)
    elif kind == 3:
        img = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    elif kind == 4:
        c = int(rng.integers(3, 24))
        cells = rng.integers(0, 2, (n, H // c + 1, W // c + 1, 1), dtype=np.uint8)
        img = (np.kron(cells, np.ones((1, c, c, 3), np.uint8))[:, :H, :W] * int(rng.integers(5, 230)) + int(rng.integers(0, 25))).astype(np.uint8)
    elif kind == 5:                                  # smooth ramps: gradients near the thresholds, many near-ties
        yy, xx = np.mgrid[0:H, 0:W]
        a, b = rng.uniform(-1.5, 1.5, 2)
        g = (128 + a * (yy - H / 2) + b * (xx - W / 2) + 20 * np.sin(xx / rng.uniform(3, 30)) * np.cos(yy / rng.uniform(3, 30)))
        img = np.repeat(np.clip(g, 0, 255).astype(np.uint8)[None, :, :, None], 3, axis=3).repeat(n, axis=0)
    else:
        img = np.full((n, H, W, 3), int(rng.integers(0, 256)), dtype=np.uint8)
        if rng.random() < 0.5:                       # one bright rectangle
            y0, x0 = int(rng.integers(0, H - 4)), int(rng.integers(0, W - 4))
            img[:, y0:y0 + int(rng.integers(2, H)), x0:x0 + int(rng.integers(2, W))] = int(rng.integers(0, 256))
    return np.ascontiguousarray(img), ("screw", "object", "texture", "noise", "blocks", "ramp", "flat")[kind]


def main():
    from oracle import objmask
    budget = float(sys.argv[1]) if len(sys.argv) > 1 else 60.0
    seed = int(sys.argv[2]) if len(sys.argv) > 2 else 0
    rng = np.random.default_rng(seed)
    t0 = time.time()
    k = nimg = bad = 0
    while time.time() - t0 < budget:
        img, kind = random_images(rng, k)
        k += 1
        got = unpack_mask_bits(obj_mask_batch(torch.from_numpy(img).cuda()).cpu().numpy(), img.shape[2])
        for i in range(len(img)):
            want = objmask.obj_mask_array(img[i])
            d = int(np.count_nonzero(got[i] != want))
            if d:
                bad += 1
                print(f"   MISMATCH [{k}] {kind} {img.shape} image {i}: {d} pixels (oracle area {int(want.sum())}, device {int(got[i].sum())})", flush=True)
        nimg += len(img)
        if k % 50 == 0:
            print(f"[{k}] {nimg} images, {bad} mismatches", flush=True)
    print(f"fuzz obj_mask: {k} batches, {nimg} images, {bad} mismatches, seed {seed}, {time.time() - t0:.0f} s")
    return 1 if bad else 0


if __name__ == "__main__":
    sys.exit(main())
