"""Per-image parameter records: host-side draws and the packed device records.

The reference draws its randomness inside ``__getitem__`` from three global RNGs
(``random``, ``np.random``, ``torch``).  Here every draw is made
up front on the host, vectorised over the batch, and handed to the CUDA path as an
explicit 192-byte record per image (``SsadImageParams`` in ``include/ssad.h``).

Two views of the same batch:

* ``draws``   -- dict of numpy arrays holding the *semantic* draws (angle, scale,
                 area fraction, aspect, polygon per-side draws, ...): what the
                 reference's own code would have drawn.  The CPU oracle consumes
                 these (``ParamBatch.image_draws(i)``).
* ``records`` -- numpy structured array (``IMAGE_PARAMS_DTYPE``) with everything
                 *derived* on the host exactly as the reference / Pillow /
                 torchvision derive it in Python float64 (rectangle size,
                 vertex list, 16.16 fixed-point affine coefficients, rotated scar
                 size): what the kernels consume.
"""
from __future__ import annotations

import math
from dataclasses import dataclass
from typing import Dict, Optional, Tuple

import numpy as np

from .config import CPP, GeneratorConfig

# Must match ``struct SsadImageParams`` (include/ssad.h) byte for byte.
IMAGE_PARAMS_DTYPE = np.dtype([
    ("warp_fix", "<i4", (6,)),       # 0
    ("scar_fix", "<i4", (6,)),       # 24
    ("centre_r", "<u4", (6,)),       # 48
    ("bright_f", "<f4", (2,)),       # 72
    ("jit_f", "<f4", (4,)),          # 80
    ("fallback_c", "<i2", (6, 2)),   # 96
    ("poly_xy", "<i2", (8, 2)),      # 120
    ("rect", "<i2", (4,)),           # 152  l, t, pw, ph (inside the donor window)
    ("scar_size", "<i2", (2,)),      # 160  nw, nh
    ("donor_off", "<i2", (2,)),      # 164  x, y of the donor window inside the donor image
    ("crop_off", "<i2", (2,)),       # 168  x, y of the working window inside src / mask
    ("donor_idx", "<i2"),            # 172  -1: the image itself
    ("label", "u1"),                 # 174
    ("warp_en", "u1"),               # 175
    ("src_variant", "u1"),           # 176
    ("scar_k", "u1"),                # 177
    ("poly_n", "u1"),                # 178
    ("rnd_rgb", "u1", (3,)),         # 179
    ("jit_order", "u1", (4,)),       # 182
    ("jit_en", "u1"),                # 186  bit0 brightness, bit1 contrast, bit2 saturation
    ("line_n", "u1"),                # 187  label 3: points of the poly-line
    ("src_index", "<u4"),            # 188  source image when the batch is indexed into a resident set
], align=False)
assert IMAGE_PARAMS_DTYPE.itemsize == 192

_DRAW_SPECS = {
    "label": ("i8", ()), "warp_angle": ("f8", ()), "warp_scale": ("f8", ()),
    "donor_idx": ("i8", ()), "crop_left": ("i8", ()), "crop_top": ("i8", ()),
    "donor_i": ("i8", ()), "donor_j": ("i8", ()), "centre_r": ("u8", (6,)),
    "variant": ("i8", ()), "area_u": ("f8", ()), "asp0": ("f8", ()), "asp1": ("f8", ()),
    "asp_pick": ("i8", ()), "rect_l": ("i8", ()), "rect_t": ("i8", ()), "rnd_rgb": ("i8", (3,)),
    "low": ("f8", ()), "high": ("f8", ()), "pick1": ("i8", ()), "pick2": ("i8", ()),
    "fallback_c": ("i8", (6, 2)), "poly_n_side": ("i8", (4,)), "poly_v": ("i8", (4, 2)),
    "scar_k": ("i8", ()), "scar_angle": ("i8", ()), "jit_order": ("i8", (4,)),
    "jit_b": ("f8", ()), "jit_c": ("f8", ()), "jit_s": ("f8", ()),
    # label 3 "long line" (datasets.py:357-388): side, the 60 (30 in patch mode) index draws, colour, split pick
    "line_side": ("i8", ()), "line_r": ("u8", (60,)), "line_rgb": ("i8", ()), "line_k": ("i8", ()),
}

LINE_COLOURS = ((0, 0, 0), (255, 255, 255), (192, 192, 192))      # PIL 'black', 'white', 'silver' (datasets.py:369)
LINE_MAX_POINTS = 32                                               # 30 points in patch mode, 6 after the split image-level


def empty_draws(batch: int) -> Dict[str, np.ndarray]:
    return {k: np.zeros((batch,) + shp, dtype=dt) for k, (dt, shp) in _DRAW_SPECS.items()}


# ------------------------------------------------------------------ derivations
def fix16(v: float) -> int:
    """Pillow ``FIX``: floor(v * 65536 + 0.5) (Geometry.c)."""
    return int(math.floor(v * 65536.0 + 0.5))


def affine_fixed_coeffs(m):
    """Six int32 coefficients of Pillow's ``affine_fixed`` for inverse matrix ``m``."""
    return (fix16(m[0]), fix16(m[1]), fix16(m[2] + m[0] * 0.5 + m[1] * 0.5),
            fix16(m[3]), fix16(m[4]), fix16(m[5] + m[3] * 0.5 + m[4] * 0.5))


def random_affine_matrix(w: int, h: int, angle: float, scale: float):
    """torchvision ``_get_inverse_affine_matrix([w/2, h/2], angle, [0,0], scale,
    [0,0])`` -- what ``RandomAffine(3, scale=(1.05,1.1))`` hands to Pillow
    (datasets.py:221-222)."""
    rot = math.radians(angle)
    cx, cy = w * 0.5, h * 0.5
    a = math.cos(rot) / 1.0
    b = -math.cos(rot) * 0.0 / 1.0 - math.sin(rot)
    c = math.sin(rot) / 1.0
    d = -math.sin(rot) * 0.0 / 1.0 + math.cos(rot)
    m = [d, -b, 0.0, -c, a, 0.0]
    m = [x / scale for x in m]
    m[2] += m[0] * (-cx) + m[1] * (-cy)
    m[5] += m[3] * (-cx) + m[4] * (-cy)
    m[2] += cx
    m[5] += cy
    return m


def rotate_expand_matrix(w: int, h: int, angle: float):
    """``Image.rotate(angle, expand=True)`` (datasets.py:344): inverse matrix and the
    expanded output size, in the same float64 operation order as PIL/Image.py."""
    angle = angle % 360.0
    if angle == 0:
        return [1.0, 0.0, 0.0, 0.0, 1.0, 0.0], w, h
    cx, cy = w / 2, h / 2
    a = -math.radians(angle)
    m = [round(math.cos(a), 15), round(math.sin(a), 15), 0.0,
         round(-math.sin(a), 15), round(math.cos(a), 15), 0.0]

    def T(x, y):
        return m[0] * x + m[1] * y + m[2], m[3] * x + m[4] * y + m[5]

    m[2], m[5] = T(-cx, -cy)
    m[2] += cx
    m[5] += cy
    xs, ys = zip(*(T(x, y) for x, y in ((0, 0), (w, 0), (w, h), (0, h))))
    nw = math.ceil(max(xs)) - math.floor(min(xs))
    nh = math.ceil(max(ys)) - math.floor(min(ys))
    m[2], m[5] = T(-(nw - w) / 2.0, -(nh - h) / 2.0)
    return m, nw, nh


def patch_size_from_draws(img_w: int, img_h: int, area_u, asp):
    """dataset_generator.py:172-180, vectorised (float64 ``np.sqrt`` as the reference)."""
    area = np.asarray(area_u, dtype=np.float64) * (img_w * img_h)
    asp = np.asarray(asp, dtype=np.float64)
    pw = np.sqrt(area * asp).astype(np.int64)
    ph = np.sqrt(area / asp).astype(np.int64)
    return np.maximum(pw, 2), np.maximum(ph, 2)


def polygon_vertices(pw: int, ph: int, n_side, v):
    """Vertex list of ``rect2poly(..., sides=8)`` (dataset_generator.py:63-98)."""
    pts = []
    for side in range(4):
        n = int(n_side[side])
        a, b = int(v[side][0]), int(v[side][1])
        if side == 0:
            pts += [(0, a)] if n == 1 else [(0, a), (0, b)]
        elif side == 1:
            pts += [(a, 0)] if n == 1 else [(a, 0), (b, 0)]
        elif side == 2:
            pts += [(pw, a)] if n == 1 else [(pw, a), (pw, b)]
        else:
            pts += [(a, ph)] if n == 1 else [(a, ph), (b, ph)]
    return pts


def fix16_array(v: np.ndarray) -> np.ndarray:
    """``fix16`` on a float64 array."""
    return np.floor(v * 65536.0 + 0.5).astype(np.int64)


def _affine_fixed_array(m0, m1, m2, m3, m4, m5) -> np.ndarray:
    """``affine_fixed_coeffs`` on arrays -> int32 [n, 6]."""
    out = np.stack([fix16_array(m0), fix16_array(m1), fix16_array(m2 + m0 * 0.5 + m1 * 0.5),
                    fix16_array(m3), fix16_array(m4), fix16_array(m5 + m3 * 0.5 + m4 * 0.5)], axis=1)
    return out.astype(np.int32)


def random_affine_fixed(w: int, h: int, angle: np.ndarray, scale: np.ndarray) -> np.ndarray:
    """``affine_fixed_coeffs(random_affine_matrix(w, h, angle[i], scale[i]))`` for a batch -> int32 [n, 6].
    (In ``random_affine_matrix`` b = -0.0 - sin = -sin and d = -0.0 + cos = cos exactly, so the inverse matrix is
    [cos, sin, 0, -sin, cos, 0] / scale before the translation terms.)"""
    angle = np.asarray(angle, dtype=np.float64)
    scale = np.asarray(scale, dtype=np.float64)
    # libm's cos / sin through math.* (NumPy's vector versions may differ in the last place); math.radians(x) is the one
    # IEEE product x * (pi / 180), so that part is an array op
    rot = (angle * (math.pi / 180.0)).tolist()
    n = len(rot)
    cs = np.fromiter(map(math.cos, rot), dtype=np.float64, count=n)
    sn = np.fromiter(map(math.sin, rot), dtype=np.float64, count=n)
    cx, cy = w * 0.5, h * 0.5
    m0, m1, m3, m4 = cs / scale, sn / scale, (-sn) / scale, cs / scale
    m2 = 0.0 / scale
    m5 = 0.0 / scale
    m2 = m2 + (m0 * (-cx) + m1 * (-cy))
    m5 = m5 + (m3 * (-cx) + m4 * (-cy))
    m2 = m2 + cx
    m5 = m5 + cy
    return _affine_fixed_array(m0, m1, m2, m3, m4, m5)


_ROT_TABLE = []


def _rot_table() -> np.ndarray:
    """[721, 4] float64: round(cos, 15), round(sin, 15), round(-sin, 15), round(cos, 15) of -radians(angle % 360)
    for the integer angles -360..360 (row angle + 360), computed with math.* as PIL/Image.py does (datasets.py:342-344)."""
    if not _ROT_TABLE:
        rows = []
        for ang in range(-360, 361):
            a = -math.radians(float(ang) % 360.0)
            rows.append((round(math.cos(a), 15), round(math.sin(a), 15), round(-math.sin(a), 15), round(math.cos(a), 15)))
        _ROT_TABLE.append(np.array(rows, dtype=np.float64))
    return _ROT_TABLE[0]


def rotate_expand_fixed(w: np.ndarray, h: np.ndarray, angle: np.ndarray):
    """``rotate_expand_matrix`` + ``affine_fixed_coeffs`` for a batch of integer angles ->
    (int32 [n, 6] coefficients, int64 [n, 2] expanded (nw, nh))."""
    w = np.asarray(w, dtype=np.float64)
    h = np.asarray(h, dtype=np.float64)
    angle = np.asarray(angle)
    ai = angle.astype(np.int64)
    if np.any(ai != angle) or np.any(np.abs(ai) > 360):
        # non-integer angles: the scalar routine (not on the reference's path, kept for callers of the function API)
        res = [rotate_expand_matrix(int(w[i]), int(h[i]), float(angle[i])) for i in range(len(w))]
        fix = np.array([affine_fixed_coeffs(r[0]) for r in res], dtype=np.int32).reshape(len(w), 6)
        return fix, np.array([(r[1], r[2]) for r in res], dtype=np.int64).reshape(len(w), 2)
    t = _rot_table()[ai + 360]
    m0, m1, m3, m4 = t[:, 0], t[:, 1], t[:, 2], t[:, 3]
    cx, cy = w / 2, h / 2
    m2 = (m0 * (-cx) + m1 * (-cy)) + 0.0
    m5 = (m3 * (-cx) + m4 * (-cy)) + 0.0
    m2 = m2 + cx
    m5 = m5 + cy
    zero = np.zeros_like(w)
    xs = np.stack([m0 * x + m1 * y + m2 for x, y in ((zero, zero), (w, zero), (w, h), (zero, h))])
    ys = np.stack([m3 * x + m4 * y + m5 for x, y in ((zero, zero), (w, zero), (w, h), (zero, h))])
    nw = np.ceil(xs.max(axis=0)) - np.floor(xs.min(axis=0))
    nh = np.ceil(ys.max(axis=0)) - np.floor(ys.min(axis=0))
    tx, ty = -(nw - w) / 2.0, -(nh - h) / 2.0
    m2, m5 = m0 * tx + m1 * ty + m2, m3 * tx + m4 * ty + m5
    fix = _affine_fixed_array(m0, m1, m2, m3, m4, m5)
    size = np.stack([nw, nh], axis=1).astype(np.int64)
    ident = (ai % 360) == 0                       # Image.rotate returns a copy: identity matrix, same size
    if np.any(ident):
        k = int(ident.sum())
        one, zer = np.ones(k), np.zeros(k)
        fix[ident] = _affine_fixed_array(one, zer, zer, zer, one, zer)
        size[ident] = np.stack([w[ident], h[ident]], axis=1).astype(np.int64)
    return fix, size


def polygon_vertices_batch(pw: np.ndarray, ph: np.ndarray, n_side: np.ndarray, v: np.ndarray):
    """``polygon_vertices`` for a batch -> (uint8 [n] vertex counts, int16 [n, 8, 2] vertices in draw order)."""
    n = len(pw)
    pw = np.asarray(pw, dtype=np.int64)
    ph = np.asarray(ph, dtype=np.int64)
    a, b = v[:, :, 0].astype(np.int64), v[:, :, 1].astype(np.int64)
    zero = np.zeros(n, dtype=np.int64)
    # candidate points, two per side: (x, y)
    cand = np.empty((n, 8, 2), dtype=np.int64)
    cand[:, 0] = np.stack([zero, a[:, 0]], 1); cand[:, 1] = np.stack([zero, b[:, 0]], 1)      # left
    cand[:, 2] = np.stack([a[:, 1], zero], 1); cand[:, 3] = np.stack([b[:, 1], zero], 1)      # top
    cand[:, 4] = np.stack([pw, a[:, 2]], 1);   cand[:, 5] = np.stack([pw, b[:, 2]], 1)        # right
    cand[:, 6] = np.stack([a[:, 3], ph], 1);   cand[:, 7] = np.stack([b[:, 3], ph], 1)        # bottom
    valid = np.ones((n, 8), dtype=bool)
    valid[:, 1::2] = np.asarray(n_side) == 2
    pos = np.cumsum(valid, axis=1) - 1
    out = np.zeros((n, 8, 2), dtype=np.int16)
    rows = np.broadcast_to(np.arange(n)[:, None], (n, 8))
    out[rows[valid], pos[valid]] = cand[valid]
    return valid.sum(axis=1).astype(np.uint8), out


# ------------------------------------------------------------------ the batch
@dataclass
class ParamBatch:
    cfg: GeneratorConfig
    draws: Dict[str, np.ndarray]
    records: np.ndarray
    line_xy: Optional[np.ndarray] = None      # int16 [B, LINE_MAX_POINTS, 2]: label-3 poly-lines once resolved against the masks

    def __len__(self) -> int:
        return self.records.shape[0]

    def image_draws(self, i: int) -> Dict:
        return {k: (v[i].copy() if v[i].ndim else v[i].item()) for k, v in self.draws.items()}

    def slice(self, lo: int, hi: int) -> "ParamBatch":
        return ParamBatch(self.cfg, {k: v[lo:hi] for k, v in self.draws.items()}, self.records[lo:hi],
                          None if self.line_xy is None else self.line_xy[lo:hi])

    def resolve_lines(self, window_mask_of) -> None:
        """Label 3: turn the index draws into poly-line vertices (``line_xy``, ``records['line_n']``).  The picks index
        ``coords_map`` (datasets.py:261-264), so they need the mask window of every label-3 image:
        ``window_mask_of(i)`` -> bool [h, w] (None: all-positive mask).  Host work: ~60 picks, a stable sort and SciPy's
        ``savgol_filter`` per image, exactly the reference's calls."""
        rec, d = self.records, self.draws
        B = len(rec)
        self.line_xy = np.zeros((B, LINE_MAX_POINTS, 2), dtype=np.int16)
        rec["line_n"] = 0
        W, H = self.cfg.work_size
        for i in np.nonzero(rec["label"] == 3)[0]:
            m = window_mask_of(int(i))
            flat = None if m is None else np.flatnonzero(np.asarray(m, dtype=bool).reshape(-1))
            n_coords = W * H if flat is None else int(flat.shape[0])
            pts = line_points(n_coords, flat, W, d["line_r"][i], int(d["line_side"][i]), int(d["line_k"][i]),
                              self.cfg.line_patch_mode)
            n = 0 if pts is None else pts.shape[0]
            if n:
                self.line_xy[i, :n] = np.clip(np.trunc(pts), -32768, 32767).astype(np.int16)   # _draw_lines: (int)p[k]
            rec["line_n"][i] = n

    def shard(self, rank: int, world: int) -> "ParamBatch":
        """This rank's contiguous block of the batch (one process per GPU; images are independent, so the shards
        never exchange anything: SURVEY.md section 8e)."""
        lo, hi = shard_bounds(len(self), rank, world)
        return self.slice(lo, hi)


def shard_bounds(batch: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block split of ``batch`` images over ``world`` ranks; the first ``batch % world`` ranks get one more."""
    if not 0 <= rank < world:
        raise ValueError("rank must be in [0, world)")
    q, r = divmod(int(batch), int(world))
    lo = rank * q + min(rank, r)
    return lo, lo + q + (1 if rank < r else 0)


def line_points(n_coords: int, flat: Optional[np.ndarray], width: int, r, side: int, k: int, patch_mode: bool):
    """datasets.py:359-377: n = 60 (30) picks ``coords_map[randint(c, int(N * i / n))]`` with c the previous index,
    (sort by x when side == 'left'), ``savgol_filter(points, 10, 2, axis=0)``, image-level: part k of
    ``np.array_split(points, 10)``.  ``flat`` = row-major flat indices of the set mask pixels (None: every pixel),
    ``r`` the 32-bit draws: randint(a, b) = a + ((r * (b - a + 1)) >> 32).  -> float64 [m, 2] (x, y), or None when the
    mask is empty (the reference raises IndexError there)."""
    from scipy.signal import savgol_filter
    if n_coords <= 0:
        return None
    n = 30 if patch_mode else 60
    c = 0
    idx = np.empty(n, dtype=np.int64)
    for i in range(n):
        hi = int(n_coords * (i / n))
        c = c + ((int(r[i]) * (hi - c + 1)) >> 32)
        idx[i] = c
    pix = idx if flat is None else flat[idx]
    pts = np.stack([pix % width, pix // width], axis=1)          # coords_map rows are (x, y)
    if side == 0:                                                # 'left': points.sort(key=lambda tup: tup[0]) -- stable
        pts = pts[np.argsort(pts[:, 0], kind="stable")]
    pts = savgol_filter([tuple(p) for p in pts], 10, 2, axis=0)
    if not patch_mode:
        pts = np.array_split(pts, 10)[k]
    return np.asarray(pts, dtype=np.float64)


def derive_records(cfg: GeneratorConfig, draws: Dict[str, np.ndarray]) -> np.ndarray:
    """draws -> packed device records (everything the reference derives in Python)."""
    B = draws["label"].shape[0]
    rec = np.zeros(B, dtype=IMAGE_PARAMS_DTYPE)
    W, H = cfg.work_size
    dW, dH = cfg.donor_window
    label = draws["label"].astype(np.int64)
    rec["label"] = label
    rec["warp_en"] = 1 if cfg.warp else 0
    rec["src_variant"] = draws["variant"]
    rec["scar_k"] = draws["scar_k"]
    rec["rnd_rgb"] = draws["rnd_rgb"]
    rec["centre_r"] = draws["centre_r"].astype(np.uint32)
    rec["fallback_c"] = draws["fallback_c"]
    rec["donor_idx"] = draws["donor_idx"] if cfg.is_texture else -1
    pc = cfg.precrop
    if cfg.patch_localization:
        rec["crop_off"][:, 0] = pc[0] + draws["crop_left"]
        rec["crop_off"][:, 1] = pc[1] + draws["crop_top"]
        rec["donor_off"][:, 0] = draws["donor_j"]
        rec["donor_off"][:, 1] = draws["donor_i"]
    # brightness factors: enhance(random.choice([low, high])) twice; Blend.c takes (float)alpha
    lh = np.stack([draws["low"], draws["high"]], axis=1)
    rec["bright_f"][:, 0] = lh[np.arange(B), draws["pick1"]].astype(np.float32)
    rec["bright_f"][:, 1] = lh[np.arange(B), draws["pick2"]].astype(np.float32)
    # jitter
    rec["jit_order"] = draws["jit_order"]
    rec["jit_f"][:, 0] = draws["jit_b"].astype(np.float32)
    rec["jit_f"][:, 1] = draws["jit_c"].astype(np.float32)
    rec["jit_f"][:, 2] = draws["jit_s"].astype(np.float32)
    rec["jit_f"][:, 3] = 0.0
    rec["jit_en"] = 0b0111 if cfg.jitter else 0
    # rectangle
    asp = np.where(draws["asp_pick"] == 0, draws["asp0"], draws["asp1"])
    pw, ph = patch_size_from_draws(dW, dH, draws["area_u"], asp)
    rec["rect"][:, 0] = draws["rect_l"]
    rec["rect"][:, 1] = draws["rect_t"]
    rec["rect"][:, 2] = pw
    rec["rect"][:, 3] = ph
    # Python-float derivations, vectorised: +, -, *, / and floor on float64 arrays round exactly as the scalar Python
    # expressions of torchvision / PIL do when written in the same order; cos / sin go through math.* (NumPy's SIMD
    # versions may differ in the last bit), once per image for the warp and from a 91-entry table for the scar angles.
    if cfg.warp:
        ang, sc = draws["warp_angle"], draws["warp_scale"]
        if np.any(ang == 0.0):
            raise ValueError("warp_angle == 0 takes Pillow's ImagingScaleAffine path, which is not modelled")
        rec["warp_fix"] = random_affine_fixed(W, H, ang, sc)
    is_poly = label == 1
    if np.any(is_poly):
        idx = np.nonzero(is_poly)[0]
        n, xy = polygon_vertices_batch(pw[idx], ph[idx], draws["poly_n_side"][idx], draws["poly_v"][idx])
        rec["poly_n"][idx] = n
        rec["poly_xy"][idx] = xy
    is_scar = label == 2
    if np.any(is_scar):
        idx = np.nonzero(is_scar)[0]
        fix, size = rotate_expand_fixed(pw[idx], ph[idx], draws["scar_angle"][idx])
        rec["scar_fix"][idx] = fix
        rec["scar_size"][idx] = size
    is_line = label == 3
    if np.any(is_line):
        # a line is a flat colour behind a coverage mask: no patch rectangle, no cut, the colour in rnd_rgb; the
        # vertices depend on the mask and are resolved later (ParamBatch.resolve_lines)
        idx = np.nonzero(is_line)[0]
        rec["rect"][idx] = 0
        rec["src_variant"][idx] = 2
        rec["rnd_rgb"][idx] = np.asarray(LINE_COLOURS, dtype=np.uint8)[draws["line_rgb"][idx]]
    return rec


def sample_params(cfg: GeneratorConfig, batch: int, rng: Optional[np.random.Generator] = None,
                  labels: Optional[np.ndarray] = None) -> ParamBatch:
    """Draw one batch of parameters with the reference's distributions
    (file:line beside each draw)."""
    rng = np.random.default_rng() if rng is None else rng
    B = int(batch)
    d = empty_draws(B)
    W, H = cfg.work_size
    dW, dH = cfg.donor_window
    iW, iH = cfg.imsize
    # datasets.py:215  y = random.randint(0, 3)  (3-class contract: {0,1,2})
    d["label"] = rng.integers(0, cfg.num_labels, B) if labels is None else np.asarray(labels, dtype=np.int64).copy()
    label = d["label"]
    if np.any(label > 3) or np.any(label < 0):
        raise ValueError("labels must be in {0,1,2,3}")
    # datasets.py:221  RandomAffine(3, scale=(1.05, 1.1)): torch draws float32
    ang = rng.uniform(-CPP.affine_degrees, CPP.affine_degrees, B).astype(np.float32)
    ang[ang == 0] = np.float32(1e-3)
    d["warp_angle"] = ang.astype(np.float64)
    d["warp_scale"] = rng.uniform(*CPP.affine_scale, B).astype(np.float32).astype(np.float64)
    # datasets.py:226  random.choice(self.images_for_cut)
    d["donor_idx"] = rng.integers(0, max(cfg.n_donors, 1), B) if cfg.is_texture else np.full(B, -1)
    if cfg.patch_localization:
        pc = cfg.precrop
        xw, xh = pc[2] - pc[0], pc[3] - pc[1]
        ps = cfg.patch_size
        if ps > xw or ps > xh:
            raise ValueError("patch_size larger than the (pre-cropped) image: the reference's randint raises here")
        d["crop_left"] = rng.integers(0, xw - ps + 1, B)      # datasets.py:249
        d["crop_top"] = rng.integers(0, xh - ps + 1, B)       # datasets.py:250
        d["donor_i"] = rng.integers(0, iH - ps + 1, B)        # RandomCrop.get_params
        d["donor_j"] = rng.integers(0, iW - ps + 1, B)
    d["centre_r"] = rng.integers(0, 1 << 32, (B, 6), dtype=np.uint64)
    d["variant"] = rng.choice(3, size=B, p=CPP.variant_probabilities)       # datasets.py:270,311
    is_scar = label == 2
    ar = np.where(is_scar[:, None], np.asarray(cfg.scar_area_ratio)[None], np.asarray(cfg.patch_area_ratio)[None])
    a0 = np.where(is_scar[:, None], np.asarray(cfg.scar_aspect_ratio[0])[None], np.asarray(cfg.rectangle_aspect_ratio[0])[None])
    a1 = np.where(is_scar[:, None], np.asarray(cfg.scar_aspect_ratio[1])[None], np.asarray(cfg.rectangle_aspect_ratio[1])[None])
    d["area_u"] = rng.uniform(ar[:, 0], ar[:, 1])                           # dataset_generator.py:173
    d["asp0"] = rng.uniform(a0[:, 0], a0[:, 1])                             # :174
    d["asp1"] = rng.uniform(a1[:, 0], a1[:, 1])
    d["asp_pick"] = rng.integers(0, 2, B)
    asp = np.where(d["asp_pick"] == 0, d["asp0"], d["asp1"])
    pw, ph = patch_size_from_draws(dW, dH, d["area_u"], asp)
    d["rect_l"] = rng.integers(0, np.maximum(dW - pw, 1) + 1)               # :183-190 (inclusive)
    d["rect_t"] = rng.integers(0, np.maximum(dH - ph, 1) + 1)
    d["rnd_rgb"] = rng.integers(0, 256, (B, 3))                             # :195-199
    d["low"] = rng.uniform(*CPP.brightness_low, B)                          # datasets.py:296-297
    d["high"] = rng.uniform(*CPP.brightness_high, B)
    d["pick1"] = rng.integers(0, 2, B)
    d["pick2"] = rng.integers(0, 2, B)
    box = np.where(is_scar[:, None], np.asarray(cfg.box_scar)[None], np.asarray(cfg.box_poly)[None])
    d["fallback_c"][:, :, 0] = rng.integers(box[:, 0:1], box[:, 2:3] + 1, (B, 6))   # dataset_generator.py:113-114
    d["fallback_c"][:, :, 1] = rng.integers(box[:, 1:2], box[:, 3:4] + 1, (B, 6))
    # rect2poly draws, dataset_generator.py:65-98
    ns = rng.integers(1, 3, (B, 4))
    d["poly_n_side"] = ns
    hh, hw = (ph / 2).astype(np.int64), (pw / 2).astype(np.int64)
    one = np.ones(B, dtype=np.int64)
    two = ns == 2
    v = np.zeros((B, 4, 2), dtype=np.int64)
    # left: n==1 -> randint(1,h); n==2 -> randint(h/2+1,h), randint(1,h/2)
    v[:, 0, 0] = np.where(two[:, 0], rng.integers(hh + 1, ph + 1), rng.integers(one, ph + 1))
    v[:, 0, 1] = rng.integers(one, hh + 1)
    # top: n==1 -> randint(1,w); n==2 -> randint(1,w/2), randint(w/2+1,w)
    v[:, 1, 0] = np.where(two[:, 1], rng.integers(one, hw + 1), rng.integers(one, pw + 1))
    v[:, 1, 1] = rng.integers(hw + 1, pw + 1)
    # right: n==1 -> randint(1,h); n==2 -> randint(1,h/2), randint(h/2+1,h)
    v[:, 2, 0] = np.where(two[:, 2], rng.integers(one, hh + 1), rng.integers(one, ph + 1))
    v[:, 2, 1] = rng.integers(hh + 1, ph + 1)
    # bottom: n==1 -> randint(1,w); n==2 -> randint(w/2+1,w), randint(1,w/2)
    v[:, 3, 0] = np.where(two[:, 3], rng.integers(hw + 1, pw + 1), rng.integers(one, pw + 1))
    v[:, 3, 1] = rng.integers(one, hw + 1)
    v[:, :, 1] = np.where(two, v[:, :, 1], 0)
    d["poly_v"] = v
    d["scar_k"] = rng.integers(2, 6, B)                                     # datasets.py:341
    d["scar_angle"] = rng.integers(-45, 46, B)                              # :342
    # ColorJitter.get_params: randperm(4); b, c, s ~ U(0.9, 1.1) drawn as float32
    d["jit_order"] = rng.permuted(np.tile(np.arange(4), (B, 1)), axis=1)
    lo, hi = 1 - CPP.jitter_offset, 1 + CPP.jitter_offset
    for k in ("jit_b", "jit_c", "jit_s"):
        d[k] = rng.uniform(lo, hi, B).astype(np.float32).astype(np.float64)
    if cfg.num_labels > 3 or np.any(label == 3):
        # label 3 (drawn last: batches of the 3-class contract keep the values they had before this label existed)
        d["line_side"] = rng.integers(0, 2, B)                                  # datasets.py:359 random.choice(['left','top'])
        d["line_r"] = rng.integers(0, 1 << 32, (B, 60), dtype=np.uint64)        # :364 random.randint(c, int(N * i / n))
        d["line_rgb"] = rng.integers(0, 3, B)                                   # :369 random.choice(['black','white','silver'])
        d["line_k"] = rng.integers(0, 10, B)                                    # :376 random.randint(0, 9)
    return ParamBatch(cfg, d, derive_records(cfg, d))


def params_from_draws(cfg: GeneratorConfig, draws: Dict[str, np.ndarray]) -> ParamBatch:
    """Build a batch from explicit draws (replay / tests)."""
    B = np.asarray(draws["label"]).shape[0]
    full = empty_draws(B)
    for k, v in draws.items():
        if k not in full:
            raise KeyError(k)
        full[k] = np.asarray(v, dtype=full[k].dtype).reshape(full[k].shape)
    return ParamBatch(cfg, full, derive_records(cfg, full))
