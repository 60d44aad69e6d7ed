"""Synthetic MVTec-like inputs (there is no dataset on the build / GPU boxes).

Images are low-frequency colour noise (so that the similarity branch of
datasets.py:295 takes both outcomes) with an optional composited object, as laid
out in SURVEY.md section 8(d)."""
from __future__ import annotations

import numpy as np


def _upsample(a: np.ndarray, h: int, w: int) -> np.ndarray:
    """separable linear upsampling of a coarse [gh, gw, c] grid to [h, w, c] (float)."""
    gh, gw = a.shape[:2]
    ys = np.linspace(0, gh - 1, h)
    xs = np.linspace(0, gw - 1, w)
    y0 = np.floor(ys).astype(int).clip(0, gh - 2)
    x0 = np.floor(xs).astype(int).clip(0, gw - 2)
    fy = (ys - y0)[:, None, None]
    fx = (xs - x0)[None, :, None]
    a = a.astype(np.float64)
    top = a[y0][:, x0] * (1 - fx) + a[y0][:, x0 + 1] * fx
    bot = a[y0 + 1][:, x0] * (1 - fx) + a[y0 + 1][:, x0 + 1] * fx
    return top * (1 - fy) + bot * fy


def texture_images(n: int, h: int, w: int, seed: int, grid: int = 8, tint=None) -> np.ndarray:
    """[n, h, w, 3] u8 smooth colour noise + fine grain."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, h, w, 3), dtype=np.uint8)
    for i in range(n):
        coarse = rng.integers(0, 256, (max(h // grid, 2), max(w // grid, 2), 3))
        if tint is not None:
            coarse = (coarse * 0.35 + np.asarray(tint)[None, None, :] * 0.65)
        img = _upsample(coarse, h, w) + rng.normal(0, 4, (h, w, 3))
        out[i] = np.clip(img, 0, 255).astype(np.uint8)
    return out


def disc_mask(h: int, w: int, frac: float = 0.4) -> np.ndarray:
    """centred blob silhouette covering ~frac of the image -> bool[h, w]."""
    yy, xx = np.mgrid[0:h, 0:w]
    r2 = frac * h * w / np.pi
    return ((yy - h / 2 + 0.5) ** 2 / 1.2 + (xx - w / 2 + 0.5) ** 2 * 1.2) <= r2


def object_images(n: int, h: int, w: int, seed: int) -> np.ndarray:
    """'bottle-like' fixed-pose object: textured blob on a light background."""
    rng = np.random.default_rng(seed)
    tex = texture_images(n, h, w, seed + 1, tint=(150, 90, 60))
    m = disc_mask(h, w)[None, :, :, None]
    bg = np.clip(rng.normal(205, 3, (n, h, w, 3)), 0, 255).astype(np.uint8)
    return np.where(m, tex, bg).astype(np.uint8)


def screw_images(n: int, h: int, w: int, seed: int) -> np.ndarray:
    """non-fixed-pose object: dark elongated rounded rectangle at a random pose on a
    light background N(190, 3^2), 2-px soft edge."""
    rng = np.random.default_rng(seed)
    out = np.empty((n, h, w, 3), dtype=np.uint8)
    yy, xx = np.mgrid[0:h, 0:w].astype(np.float32)
    for i in range(n):
        ang = rng.uniform(0, np.pi)
        cx = rng.uniform(0.35, 0.65) * w
        cy = rng.uniform(0.35, 0.65) * h
        L, T = 0.32 * w, 0.06 * h
        u = (xx - cx) * np.cos(ang) + (yy - cy) * np.sin(ang)
        v = -(xx - cx) * np.sin(ang) + (yy - cy) * np.cos(ang)
        dx = np.maximum(np.abs(u) - (L - T), 0)
        dist = np.sqrt(dx * dx + v * v) - T           # signed distance to the capsule
        a = np.clip(0.5 - dist / 2.0, 0, 1)[..., None]
        obj = rng.uniform(60, 120) + 12 * np.sin(u / 3.0)[..., None] + rng.normal(0, 3, (h, w, 3))
        bg = rng.normal(190, 3, (h, w, 3))
        out[i] = np.clip(obj * a + bg * (1 - a), 0, 255).astype(np.uint8)
    return out
