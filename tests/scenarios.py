"""Seeded scenarios shared by the golden-vector generator, the CPU tests and the GPU
parity tests: one per mode of the reference's generator (at
sizes the CPU oracle finishes in seconds)."""
from __future__ import annotations

import os
import sys
from dataclasses import dataclass
from typing import List, Optional

import numpy as np

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
if ROOT not in sys.path:
    sys.path.insert(0, ROOT)

import ssad_b200  # noqa: E402
from ssad_b200 import GeneratorConfig, sample_params  # noqa: E402
from ssad_b200 import synth  # noqa: E402


@dataclass
class Scenario:
    name: str
    cfg: GeneratorConfig
    images: np.ndarray                 # [B, H, W, 3] u8 originals at imsize
    mask: Optional[np.ndarray]         # bool [H, W] fixed mask, or [B, H, W] per image, None = texture
    donors: Optional[np.ndarray]       # [D, H, W, 3] u8 donor bank (textures)
    params: "ssad_b200.ParamBatch"


def _force_some(labels: np.ndarray) -> np.ndarray:
    labels = labels.copy()
    labels[:3] = (0, 1, 2)
    return labels


def build(name: str, batch: int = 12, seed: int = 0) -> Scenario:
    rng = np.random.default_rng(1000 + seed)
    labels = _force_some(rng.integers(0, 3, batch))
    if name == "image_level":          # config 1: fixed-pose object, cached mask, warp on
        S = 128
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
        imgs = synth.object_images(batch, S, S, 11 + seed)
        mask = synth.disc_mask(S, S)
        donors = None
    elif name == "image_level_256":
        S = 256
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
        imgs = synth.object_images(batch, S, S, 12 + seed)
        mask = synth.disc_mask(S, S)
        donors = None
    elif name == "patch_level":        # config 2: window of a bigger image, BASELINE ratios
        S, ps = 160, 96
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S), patch_localization=True, patch_size=ps,
                              patch_area_ratio=(0.1, 0.2), scar_area_ratio=(0.01, 0.02))
        imgs = synth.object_images(batch, S, S, 13 + seed)
        mask = synth.disc_mask(S, S, 0.45)
        donors = None
    elif name == "patch_level_32":     # tools.training default: ps=32, code ratios (patch > window happens)
        S, ps = 128, 32
        cfg = GeneratorConfig(subject="capsule", imsize=(256, 256), patch_localization=True, patch_size=ps)
        S = 256
        imgs = synth.object_images(batch, S, S, 14 + seed)
        mask = synth.disc_mask(S, S, 0.35)
        donors = None
    elif name == "texture":            # config 3: all-positive mask, donor bank
        S = 128
        cfg = GeneratorConfig(subject="carpet", imsize=(S, S), n_donors=15)
        imgs = synth.texture_images(batch, S, S, 15 + seed, tint=(120, 110, 100))
        donors = np.concatenate([synth.texture_images(1, S, S, 100 + k, tint=rng.integers(40, 220, 3))
                                 for k in range(15)])
        mask = None
    elif name == "non_fixed":          # config 4 shape: per-image mask, warp off
        S = 128
        cfg = GeneratorConfig(subject="screw", imsize=(S, S))
        imgs = synth.screw_images(batch, S, S, 16 + seed)
        mask = "per_image"
        donors = None
    elif name == "empty_mask":         # edge: N == 0 -> fallback centre
        S = 96
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
        imgs = synth.object_images(batch, S, S, 17 + seed)
        mask = np.zeros((S, S), dtype=bool)
        donors = None
    elif name == "single_pixel_mask":  # edge: N == 1 -> element 0, no draw
        S = 96
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
        imgs = synth.object_images(batch, S, S, 18 + seed)
        mask = np.zeros((S, S), dtype=bool)
        mask[40, 57] = True
        donors = None
    elif name == "flat_grey":          # similarity branch always fires (cos == 1)
        S = 96
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
        imgs = np.clip(rng.normal(128, 20, (batch, S, S, 1)).repeat(3, axis=3), 0, 255).astype(np.uint8)
        mask = synth.disc_mask(S, S)
        donors = None
    elif name == "big_fixed_768":      # rows stream through chunk buffers (does not fit shared memory), warp gather
        S = 768
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S))
        imgs = synth.object_images(batch, S, S, 19 + seed)
        mask = synth.disc_mask(S, S)
        donors = None
    elif name == "big_non_fixed_1024":  # config 4 shape: 1024x1024, per-image obj_mask, no warp -> TMA streaming
        S = 1024
        cfg = GeneratorConfig(subject="screw", imsize=(S, S))
        imgs = synth.screw_images(batch, S, S, 20 + seed)
        mask = "per_image"
        donors = None
    elif name == "mid_texture_512":     # 512x512: eight CTAs per image, rows resident
        S = 512
        cfg = GeneratorConfig(subject="carpet", imsize=(S, S), n_donors=15)
        imgs = synth.texture_images(batch, S, S, 21 + seed, tint=(120, 110, 100))
        donors = np.concatenate([synth.texture_images(1, S, S, 300 + k, tint=rng.integers(40, 220, 3))
                                 for k in range(15)])
        mask = None
    elif name == "lines_image":         # label 3 "long line" (datasets.py:357-388), image-level: width 3, split in 10
        S = 128
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S), num_labels=4)
        imgs = synth.object_images(batch, S, S, 22 + seed)
        mask = synth.disc_mask(S, S)
        donors = None
        labels = np.where(np.arange(batch) % 3 != 1, 3, rng.integers(0, 4, batch))
    elif name == "lines_patch":         # label 3 in patch mode: width 1, all 30 points; coverage rule may zero the label
        S, ps = 160, 96
        cfg = GeneratorConfig(subject="bottle", imsize=(S, S), patch_localization=True, patch_size=ps, num_labels=4,
                              patch_area_ratio=(0.1, 0.2), scar_area_ratio=(0.01, 0.02))
        imgs = synth.object_images(batch, S, S, 23 + seed)
        mask = synth.disc_mask(S, S, 0.45)
        donors = None
        labels = np.where(np.arange(batch) % 4 != 2, 3, rng.integers(0, 4, batch))
    elif name == "lines_texture":       # label 3 on a texture class: all-positive mask, warp on
        S = 128
        cfg = GeneratorConfig(subject="carpet", imsize=(S, S), n_donors=15, num_labels=4)
        imgs = synth.texture_images(batch, S, S, 24 + seed, tint=(120, 110, 100))
        donors = np.concatenate([synth.texture_images(1, S, S, 400 + k, tint=rng.integers(40, 220, 3))
                                 for k in range(15)])
        mask = None
        labels = np.where(np.arange(batch) % 2 == 0, 3, rng.integers(0, 4, batch))
    elif name == "screw_patch":         # non-fixed class in patch mode: per-image mask AND the class pre-crop
        S, ps = 256, 64                 # (25, 25, 230, 230), datasets.py:246-248; all four labels
        cfg = GeneratorConfig(subject="screw", imsize=(S, S), patch_localization=True, patch_size=ps, num_labels=4)
        imgs = synth.screw_images(batch, S, S, 25 + seed)
        mask = "per_image"
        donors = None
        labels = np.arange(batch) % 4
    else:
        raise KeyError(name)
    if isinstance(mask, str):
        sys.path.insert(0, ROOT)
        from oracle import objmask
        mask = np.stack([objmask.obj_mask_array(im) for im in imgs])
    params = sample_params(cfg, batch, np.random.default_rng(2000 + seed), labels=labels)
    return Scenario(name, cfg, imgs, mask, donors, params)


BIG = ["big_fixed_768", "big_non_fixed_1024", "mid_texture_512"]      # no golden vectors (size); oracle parity only

ALL = ["image_level", "image_level_256", "patch_level", "patch_level_32", "texture", "non_fixed",
       "empty_mask", "single_pixel_mask", "flat_grey",
       "lines_image", "lines_patch", "lines_texture", "screw_patch"]


# ------------------------------------------------------------------ oracle driver
def mask_to_pil(mask: np.ndarray):
    from PIL import Image
    return Image.fromarray(mask).convert("RGB")


def run_oracle(sc: Scenario, i: int):
    """Explicit-parameter oracle for image ``i`` -> (x u8 [h,w,3], y, info)."""
    from PIL import Image
    from oracle import generator as og
    original = Image.fromarray(sc.images[i])
    donors = [Image.fromarray(d) for d in sc.donors] if sc.donors is not None else None
    if sc.mask is None:
        seg = Image.new(size=sc.cfg.imsize, mode="RGB", color="white")      # datasets.py:198
        mask_fn = None
    elif sc.mask.ndim == 3:
        seg = None
        mask_fn = lambda _orig, m=sc.mask[i]: mask_to_pil(m)               # noqa: E731
    else:
        seg = mask_to_pil(sc.mask)
        mask_fn = None
    x, y, info = og.generate_item(original, sc.params.image_draws(i), sc.cfg, donors, seg, mask_fn)
    return np.array(x), y, info


def window_mask(sc: Scenario, i: int) -> np.ndarray:
    """The mask window ``__getitem__`` builds ``coords_map`` from (bool [h, w])."""
    W, H = sc.cfg.imsize
    if sc.mask is None:
        m = np.ones((H, W), dtype=bool)
    elif sc.mask.ndim == 3:
        m = sc.mask[i]
    else:
        m = sc.mask
    if sc.cfg.patch_localization:
        pc = sc.cfg.precrop
        d = sc.params.draws
        ps = sc.cfg.patch_size
        l, t = pc[0] + int(d["crop_left"][i]), pc[1] + int(d["crop_top"][i])
        m = m[t:t + ps, l:l + ps]
    return m
