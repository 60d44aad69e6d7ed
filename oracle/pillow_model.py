"""numpy restatement of the Pillow / torchvision arithmetic on the generator path.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

The reference (``/root/reference/src/self_supervised``) is pure Python; every
pixel operation of its hot path executes inside third-party code that is NOT part
of the reference tree and is not version-pinned by it:

* Pillow (installed here: 12.2.0) -- ``Image.paste`` (``Paste.c``),
  ``Image.transform(AFFINE, NEAREST)`` (``Geometry.c: affine_fixed``),
  ``ImageDraw.polygon`` (``Draw.c: polygon_generic``), ``Image.blend``
  (``Blend.c``), ``Image.convert('L')`` (``Convert.c``), ``ImageEnhance``.
* torchvision (installed: 0.26.0) -- ``ColorJitter``, ``RandomAffine``.

Each function below restates the published algorithm of one of those routines
and cites the reference call site it serves.  ``tests/test_pillow_model.py``
checks every one of them bit-exact against the installed libraries, which is
what lets the CUDA kernels (written from these models) be compared bit-for-bit.

All float32 arithmetic is written as separate, individually rounded numpy
float32 operations (the x86-64 Pillow wheel has no FMA contraction).
"""
from __future__ import annotations

import math
from typing import List, Sequence, Tuple

import numpy as np

f32 = np.float32


# --------------------------------------------------------------------------
# A.1  Image.paste(patch, (l, t), mask)   -- dataset_generator.py:268-275
# --------------------------------------------------------------------------
def paste_select(dst: np.ndarray, src: np.ndarray, left: int, top: int,
                 alpha: np.ndarray) -> np.ndarray:
    """``dst.copy().paste(src, (left, top), mask)`` for masks holding only 0/255.

    dst [H,W,3] u8, src [ph,pw,>=3] u8, alpha [ph,pw] (bool or 0/255).  The box is
    clipped to the destination; source/mask are indexed with the same offset.
    """
    out = dst.copy()
    H, W = dst.shape[:2]
    ph, pw = src.shape[:2]
    x0, y0 = max(left, 0), max(top, 0)
    x1, y1 = min(left + pw, W), min(top + ph, H)
    if x1 <= x0 or y1 <= y0:
        return out
    s = src[y0 - top:y1 - top, x0 - left:x1 - left, :3]
    a = np.asarray(alpha)[y0 - top:y1 - top, x0 - left:x1 - left] != 0
    region = out[y0:y1, x0:x1]
    region[a] = s[a]
    return out


# --------------------------------------------------------------------------
# A.2  nearest-neighbour affine in 16.16 fixed point (Geometry.c affine_fixed)
#      datasets.py:221-222 (RandomAffine) and datasets.py:344 (scar.rotate)
# --------------------------------------------------------------------------
def fix16(v: float) -> int:
    """Pillow's ``FIX(v) = floor(v * 65536 + 0.5)``."""
    return int(math.floor(v * 65536.0 + 0.5))


def affine_fixed_coeffs(m: Sequence[float]) -> Tuple[int, int, int, int, int, int]:
    """The six int32 coefficients ``affine_fixed`` iterates with (half-pixel term
    folded into a2/a5)."""
    a0, a1, a3, a4 = fix16(m[0]), fix16(m[1]), fix16(m[3]), fix16(m[4])
    a2 = fix16(m[2] + m[0] * 0.5 + m[1] * 0.5)
    a5 = fix16(m[5] + m[3] * 0.5 + m[4] * 0.5)
    return a0, a1, a2, a3, a4, a5


def affine_nearest_fixed(src: np.ndarray, coeffs: Sequence[int], out_w: int,
                         out_h: int) -> np.ndarray:
    """Apply the fixed-point inverse map; outside pixels stay 0 (all bands)."""
    a0, a1, a2, a3, a4, a5 = (int(c) for c in coeffs)
    h, w = src.shape[:2]
    ys, xs = np.mgrid[0:out_h, 0:out_w].astype(np.int64)
    xx = a2 + ys * a1 + xs * a0
    yy = a5 + ys * a4 + xs * a3
    # int32 wrap as in C is irrelevant at these magnitudes; arithmetic shift
    xin = xx >> 16
    yin = yy >> 16
    ok = (xin >= 0) & (xin < w) & (yin >= 0) & (yin < h)
    out = np.zeros((out_h, out_w) + src.shape[2:], dtype=src.dtype)
    out[ok] = src[yin[ok], xin[ok]]
    return out


def rotate_expand_matrix(w: int, h: int, angle: float):
    """``Image.rotate(angle, expand=True)`` matrix and output size
    (site-packages/PIL/Image.py rotate()).  Returns (matrix6, nw, nh)."""
    angle = angle % 360.0
    if angle == 0:
        return [1.0, 0.0, 0.0, 0.0, 1.0, 0.0], w, h
    center = (w / 2, h / 2)
    a = -math.radians(angle)
    m = [round(math.cos(a), 15), round(math.sin(a), 15), 0.0,
         round(-math.sin(a), 15), round(math.cos(a), 15), 0.0]

    def T(x, y, mm):
        return mm[0] * x + mm[1] * y + mm[2], mm[3] * x + mm[4] * y + mm[5]

    m[2], m[5] = T(-center[0], -center[1], m)
    m[2] += center[0]
    m[5] += center[1]
    xx, yy = [], []
    for x, y in ((0, 0), (w, 0), (w, h), (0, h)):
        tx, ty = T(x, y, m)
        xx.append(tx)
        yy.append(ty)
    nw = math.ceil(max(xx)) - math.floor(min(xx))
    nh = math.ceil(max(yy)) - math.floor(min(yy))
    m[2], m[5] = T(-(nw - w) / 2.0, -(nh - h) / 2.0, m)
    return m, nw, nh


def random_affine_matrix(w: int, h: int, angle: float, scale: float) -> List[float]:
    """torchvision ``_get_inverse_affine_matrix(center=[w/2,h/2], angle,
    [0,0], scale, [0,0])`` (transforms/functional.py) as used by
    ``RandomAffine(3, scale=(1.05, 1.1))`` -- datasets.py:221."""
    rot = math.radians(angle)
    cx, cy = w * 0.5, h * 0.5
    a = math.cos(rot - 0.0) / math.cos(0.0)
    b = -math.cos(rot - 0.0) * math.tan(0.0) / math.cos(0.0) - math.sin(rot)
    c = math.sin(rot - 0.0) / math.cos(0.0)
    d = -math.sin(rot - 0.0) * math.tan(0.0) / math.cos(0.0) + math.cos(rot)
    m = [d, -b, 0.0, -c, a, 0.0]
    m = [x / scale for x in m]
    m[2] += m[0] * (-cx - 0) + m[1] * (-cy - 0)
    m[5] += m[3] * (-cx - 0) + m[4] * (-cy - 0)
    m[2] += cx
    m[5] += cy
    return m


# --------------------------------------------------------------------------
# A.3  ImageDraw.polygon(points, fill) -- Draw.c polygon_generic
#      dataset_generator.py:99-100
# --------------------------------------------------------------------------
def _round_up(f) -> int:
    f = f32(f)
    return int(math.floor(f32(f + f32(0.5)))) if f >= 0 else -int(math.floor(f32(abs(f) + f32(0.5))))


def _round_down(f) -> int:
    f = f32(f)
    return int(math.ceil(f32(f - f32(0.5)))) if f >= 0 else -int(math.ceil(f32(abs(f) - f32(0.5))))


class _Edge:
    __slots__ = ("xmin", "xmax", "ymin", "ymax", "dx", "x0", "y0")

    def __init__(self, x0, y0, x1, y1):
        self.xmin, self.xmax = (x0, x1) if x0 <= x1 else (x1, x0)
        self.ymin, self.ymax = (y0, y1) if y0 <= y1 else (y1, y0)
        if y0 == y1:
            self.dx = f32(0.0)
        else:
            self.dx = f32(f32(x1 - x0) / f32(y1 - y0))
        self.x0, self.y0 = x0, y0

    def x_at(self, y: int):
        return f32(f32(f32(y - self.y0) * self.dx) + f32(self.x0))


def polygon_edges(points: Sequence[Tuple[int, int]]):
    pts = [(int(x), int(y)) for x, y in points]
    edges = [_Edge(*pts[i], *pts[i + 1]) for i in range(len(pts) - 1)]
    if pts[-1] != pts[0]:
        edges.append(_Edge(*pts[-1], *pts[0]))
    return edges


def polygon_mask(points: Sequence[Tuple[int, int]], w: int, h: int) -> np.ndarray:
    """Filled polygon (no anti-aliasing) on a ``w x h`` canvas -> bool[h, w]."""
    out = np.zeros((h, w), dtype=bool)

    def hline(x0, y, x1):
        if 0 <= y < h:
            if x0 < 0:
                x0 = 0
            elif x0 >= w:
                return
            if x1 < 0:
                return
            elif x1 >= w:
                x1 = w - 1
            if x0 <= x1:
                out[y, x0:x1 + 1] = True

    edges = polygon_edges(points)
    if not edges:
        return out
    table = []
    ymin, ymax = h - 1, 0
    for e in edges:
        ymin = min(ymin, e.ymin)
        ymax = max(ymax, e.ymax)
        if e.ymin == e.ymax:
            hline(e.xmin, e.ymin, e.xmax)
            continue
        table.append(e)
    ymin = max(ymin, 0)
    ymax = min(ymax, h)
    for y in range(ymin, ymax + 1):
        xx: List[np.float32] = []
        for i, cur in enumerate(table):
            if cur.ymin <= y <= cur.ymax:
                xx.append(cur.x_at(y))
                if y == cur.ymax and y < ymax:
                    xx.append(xx[-1])
                elif cur.dx != 0 and len(xx) % 2 == 1 and f32(np.round(xx[-1])) == xx[-1]:
                    # np.round is half-to-even, roundf is half-away; both agree on
                    # "is this value an integer", which is all that is asked here
                    for k in range(i):
                        other = table[k]
                        if (cur.dx > 0 and other.dx <= 0) or (cur.dx < 0 and other.dx >= 0):
                            continue
                        if xx[-1] == other.x_at(y):
                            off = -1 if y == ymax else 1
                            ax = cur.x_at(y + off)
                            ao = other.x_at(y + off)
                            if k < len(xx):
                                if y == cur.ymax:
                                    xx[k] = f32(max(ax, ao) + f32(1)) if cur.dx > 0 else f32(min(ax, ao) - f32(1))
                                else:
                                    xx[k] = f32(min(ax, ao)) if cur.dx > 0 else f32(max(ax, ao) + f32(1))
                            break
        xx.sort()
        x_pos = -1 if len(xx) == 0 else 0
        for i in range(1, len(xx), 2):
            x_end = _round_down(xx[i])
            if x_end < x_pos:
                continue
            x_start = _round_up(xx[i - 1])
            if x_pos > x_start:
                x_start = x_pos
                if x_end < x_start:
                    continue
            hline(x_start, y, x_end)
            x_pos = x_end + 1
    return out


# --------------------------------------------------------------------------
# A.3b  ImageDraw.line(points, fill, width) -- Draw.c ImagingDrawLine / ImagingDrawWideLine
#       datasets.py:383-388 (label 3 "long line": width 3 image-level, 1 in patch mode)
# --------------------------------------------------------------------------
def line_segment_pixels(x0: int, y0: int, x1: int, y1: int):
    """Draw.c line8/line32: the pixels of one width-1 segment, in drawing order.  The end point itself is not drawn
    (the next segment of a poly-line starts there; ``_draw_lines`` draws the last point of the poly-line itself)."""
    out = []
    dx = x1 - x0
    xs = 1
    if dx < 0:
        dx, xs = -dx, -1
    dy = y1 - y0
    ys = 1
    if dy < 0:
        dy, ys = -dy, -1
    if dx == 0:
        for _ in range(dy):
            out.append((x0, y0)); y0 += ys
    elif dy == 0:
        for _ in range(dx):
            out.append((x0, y0)); x0 += xs
    elif dx > dy:
        n = dx
        dy += dy
        e = dy - dx
        dx += dx
        for _ in range(n):
            out.append((x0, y0))
            if e >= 0:
                y0 += ys
                e -= dx
            e += dy
            x0 += xs
    else:
        n = dy
        dx += dx
        e = dx - dy
        dy += dy
        for _ in range(n):
            out.append((x0, y0))
            if e >= 0:
                x0 += xs
                e -= dy
            e += dx
            y0 += ys
    return out


def wide_line_quad(x0: int, y0: int, x1: int, y1: int, width: int):
    """Draw.c ImagingDrawWideLine: the four vertices of the polygon drawn for a segment of ``width`` > 1 (None for a
    zero-length segment, which draws the single point)."""
    dx, dy = x1 - x0, y1 - y0
    if dx == 0 and dy == 0:
        return None
    big = math.hypot(dx, dy)
    small = (width - 1) / 2.0
    ratio_max = _round_up_d(small) / big
    ratio_min = _round_down_d(small) / big
    dxmin = _round_down_d(ratio_min * dy)
    dxmax = _round_down_d(ratio_max * dy)
    dymin = _round_up_d(ratio_min * dx)
    dymax = _round_up_d(ratio_max * dx)
    return [(x0 - dxmin, y0 + dymax), (x1 - dxmin, y1 + dymax), (x1 + dxmax, y1 - dymin), (x0 + dxmax, y0 - dymin)]


def _round_up_d(f: float) -> int:          # Draw.c ROUND_UP on a double
    return int(math.floor(f + 0.5)) if f >= 0.0 else -int(math.floor(abs(f) + 0.5))


def _round_down_d(f: float) -> int:        # Draw.c ROUND_DOWN on a double
    return int(math.ceil(f - 0.5)) if f >= 0.0 else -int(math.ceil(abs(f) - 0.5))


def line_mask(points: Sequence[Tuple[float, float]], width: int, w: int, h: int) -> np.ndarray:
    """``ImageDraw.line(points, fill, width)`` (joint=None) on a ``w x h`` canvas -> bool[h, w] of the pixels set.
    Coordinates are truncated to int as ``_draw_lines`` does (``(int)p[0]`` ...)."""
    out = np.zeros((h, w), dtype=bool)
    pts = [(int(x), int(y)) for x, y in points]            # C cast: toward zero
    for (x0, y0), (x1, y1) in zip(pts[:-1], pts[1:]):
        if width <= 1:
            for x, y in line_segment_pixels(x0, y0, x1, y1):
                if 0 <= x < w and 0 <= y < h:
                    out[y, x] = True
        else:
            quad = wide_line_quad(x0, y0, x1, y1, width)
            if quad is None:
                if 0 <= x0 < w and 0 <= y0 < h:
                    out[y0, x0] = True
            else:
                out |= polygon_mask(quad, w, h)
    if width <= 1 and len(pts) >= 2:                        # _imaging.c _draw_lines: "draw last point"
        x, y = pts[-1]
        if 0 <= x < w and 0 <= y < h:
            out[y, x] = True
    return out


# --------------------------------------------------------------------------
# A.4  colour ops: Blend.c, Convert.c (RGB->L), ImageEnhance, ColorJitter
#      datasets.py:44-47,298-299,338-339,391
# --------------------------------------------------------------------------
def rgb_to_l(img: np.ndarray) -> np.ndarray:
    """``Image.convert('L')`` for RGB u8: ITU-R 601-2 in 16.16 integer."""
    r = img[..., 0].astype(np.uint32)
    g = img[..., 1].astype(np.uint32)
    b = img[..., 2].astype(np.uint32)
    return ((19595 * r + 38470 * g + 7471 * b + 0x8000) >> 16).astype(np.uint8)


def blend_u8(A: np.ndarray, B: np.ndarray, alpha: float) -> np.ndarray:
    """``Image.blend(A, B, alpha)`` on u8 arrays (broadcastable)."""
    al = f32(alpha)
    Af = np.asarray(A).astype(np.int32)
    Bf = np.asarray(B).astype(np.int32)
    d = (Bf - Af).astype(np.float32)
    t = (Af.astype(np.float32) + (al * d).astype(np.float32)).astype(np.float32)
    if 0.0 <= float(al) <= 1.0:
        return np.trunc(t).astype(np.int64).astype(np.uint8)
    t = np.where(t <= 0, f32(0), np.where(t >= 255, f32(255), np.trunc(t)))
    return t.astype(np.uint8)


def brightness(img: np.ndarray, factor: float) -> np.ndarray:
    """``ImageEnhance.Brightness(img).enhance(factor)``."""
    return blend_u8(np.zeros_like(img), img, factor)


def contrast_mean(img: np.ndarray) -> int:
    """``int(ImageStat.Stat(img.convert('L')).mean[0] + 0.5)``."""
    L = rgb_to_l(img)
    return int(float(int(L.astype(np.int64).sum())) / float(L.size) + 0.5)


def contrast(img: np.ndarray, factor: float) -> np.ndarray:
    """``ImageEnhance.Contrast(img).enhance(factor)``."""
    m = contrast_mean(img)
    return blend_u8(np.full_like(img, m), img, factor)


def saturation(img: np.ndarray, factor: float) -> np.ndarray:
    """``ImageEnhance.Color(img).enhance(factor)``."""
    L = rgb_to_l(img)
    return blend_u8(np.repeat(L[..., None], 3, axis=-1), img, factor)


def color_jitter(img: np.ndarray, order: Sequence[int], factors: Sequence[float],
                 enabled=(True, True, True, False)) -> np.ndarray:
    """torchvision ``ColorJitter.forward`` on a PIL image with explicit
    ``fn_idx`` / factors (slot 3 = hue, off in the reference)."""
    x = img
    for fn in order:
        fn = int(fn)
        if fn == 0 and enabled[0]:
            x = brightness(x, factors[0])
        elif fn == 1 and enabled[1]:
            x = contrast(x, factors[1])
        elif fn == 2 and enabled[2]:
            x = saturation(x, factors[2])
        elif fn == 3 and enabled[3]:
            raise NotImplementedError("hue slot is disabled in the reference (datasets.py:44-47)")
    return x


# --------------------------------------------------------------------------
# A.5  Container clamp -- dataset_generator.py:15-24,104-144
# --------------------------------------------------------------------------
def container_box(imsize0: int, factor: float) -> Tuple[int, int, int, int]:
    c = int(imsize0 / 2)
    lo = int(c - (c / factor))
    hi = int(c + (c / factor))
    return lo, lo, hi, hi


def clamp_to_container(box, size, centre) -> Tuple[int, int]:
    L, T, R, Bm = box
    pw, ph = size
    cx, cy = centre
    pl = cx - int(pw / 2)
    pt = cy - int(ph / 2)
    pr = cx + int(pw / 2)
    pb = cy + int(ph / 2)
    if pr > R:
        pl = R - pw
    if pb > Bm:
        pt = Bm - ph
    if pl < L:
        pl = L
    if pt < T:
        pt = T
    return pl, pt


# --------------------------------------------------------------------------
# A.8  similarity branch -- dataset_generator.py:147-159, datasets.py:295-299
# --------------------------------------------------------------------------
def cosine_of_means(sum1, n1, sum2, n2) -> float:
    """cosine similarity of mean RGB (each /255) from exact integer sums, in the
    operation order of ``check_color_similarity`` + sklearn ``cosine_similarity``
    (normalise each row by its L2 norm, then dot)."""
    m1 = [float(np.float64(s) / np.float64(n1)) / 255 for s in sum1]
    m2 = [float(np.float64(s) / np.float64(n2)) / 255 for s in sum2]
    n_1 = math.sqrt(m1[0] * m1[0] + m1[1] * m1[1] + m1[2] * m1[2])
    n_2 = math.sqrt(m2[0] * m2[0] + m2[1] * m2[1] + m2[2] * m2[2])
    if n_1 == 0.0:
        n_1 = 1.0
    if n_2 == 0.0:
        n_2 = 1.0
    u = [v / n_1 for v in m1]
    v = [v / n_2 for v in m2]
    return u[0] * v[0] + u[1] * v[1] + u[2] * v[2]
