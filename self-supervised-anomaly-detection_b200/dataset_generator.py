"""Function-level mirror of the reference's generator module and of the exerciser script's entry points:
``src/self_supervised/dataset_generator.py`` (obj_mask, rect2poly, check_valid_coordinates_by_container,
get_random_coordinate, paste_patch, generate_patch, Container) and ``src/test_artificial_transformations.py``
(do_patch, do_scar).

Same names and argument meaning; the pixel work runs in libssad_gen.so on a CUDA device (no CPU fallback).  Where
the reference draws random numbers inside the function, these take the draws explicitly (``params=`` / ``points=``)
or sample them with the reference's distributions from an ``rng`` -- so the reference applied to the same
parameters is the oracle (tests/test_gpu_host_api.py).  PIL in -> PIL out, like the reference; tensors stay tensors.
"""
from __future__ import annotations

import ctypes as C
from typing import Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as N
from .config import CPP, GeneratorConfig
from .generator import BatchGenerator, obj_mask_batch, unpack_mask_bits
from .params import ParamBatch, patch_size_from_draws, polygon_vertices, sample_params


def _device(device=None) -> torch.device:
    if not torch.cuda.is_available():
        raise RuntimeError("a CUDA device is required: this package has no CPU fallback")
    return torch.device(device) if device is not None else torch.device("cuda", torch.cuda.current_device())


def _to_u8_hwc(image) -> Tuple[np.ndarray, bool]:
    """PIL image / ndarray -> contiguous uint8 [H, W, 3]; second value: input was a PIL image."""
    is_pil = hasattr(image, "convert")
    a = np.array(image.convert("RGB")) if is_pil else np.asarray(image)
    if a.dtype != np.uint8 or a.ndim != 3 or a.shape[2] != 3:
        raise ValueError("expected an RGB uint8 image [H, W, 3]")
    return np.ascontiguousarray(a), is_pil


def _from_u8_hwc(a: np.ndarray, as_pil: bool):
    if as_pil:
        from PIL import Image
        return Image.fromarray(a)
    return a


def _mask_bool(segmentation) -> Optional[np.ndarray]:
    """The reference's ``np.array(segmentation.convert('1'))`` (datasets.py:263): any channel set -> True."""
    if segmentation is None:
        return None
    if hasattr(segmentation, "convert"):
        return np.array(segmentation.convert("1")).astype(bool)
    m = np.asarray(segmentation)
    if m.ndim == 3:
        m = m[..., 0]
    return m.astype(bool)


# ---------------------------------------------------------------------- Container / clamp (host integer code)
class Container:
    """dataset_generator.py:15-24 -- the centred square a defect must stay in (uses ``imsize[0]`` only)."""

    def __init__(self, imsize: tuple, scaling_factor: float) -> None:
        half = int(imsize[0] / 2)
        self.center = half
        self.dim = int(imsize[0] / scaling_factor)
        self.left = self.top = int(half - (half / scaling_factor))
        self.right = self.bottom = int(half + (half / scaling_factor))
        self.width = self.right - self.left
        self.height = self.bottom - self.top


def check_valid_coordinates_by_container(imsize: tuple, patchsize: tuple, current_coords: Optional[tuple] = None,
                                         container_scaling_factor: float = 1, rng: Optional[np.random.Generator] = None):
    """dataset_generator.py:104-144: centre -> paste (left, top) kept inside the container.  The extents tested are
    ``2 * int(size / 2)`` while the corrections use the full size, and the left/top clamp runs last.
    ``current_coords=None`` draws a centre inside the container (``rng`` instead of the global ``random``)."""
    box = Container(imsize, container_scaling_factor)
    w, h = int(patchsize[0]), int(patchsize[1])
    if current_coords is None:
        rng = np.random.default_rng() if rng is None else rng
        cx, cy = int(rng.integers(box.left, box.right + 1)), int(rng.integers(box.top, box.bottom + 1))
    else:
        cx, cy = int(current_coords[0]), int(current_coords[1])
    left, top = cx - int(w / 2), cy - int(h / 2)
    if cx + int(w / 2) > box.right:
        left = box.right - w
    if cy + int(h / 2) > box.bottom:
        top = box.bottom - h
    left = max(left, box.left)
    top = max(top, box.top)
    return (left, top)


def get_random_coordinate(xy_coords, rng: Optional[np.random.Generator] = None):
    """dataset_generator.py:259-265: None for an empty list, element 0 for one element, else a uniform pick."""
    n = len(xy_coords)
    if n == 0:
        return None
    if n < 2:
        return xy_coords[0]
    rng = np.random.default_rng() if rng is None else rng
    return xy_coords[int(rng.integers(0, n))]


# ---------------------------------------------------------------------- obj_mask
def obj_mask(image, device=None):
    """dataset_generator.py:27-39 on the device: convert('L') -> canny(1.5, 5, 15) -> dilate 3x3 -> close 3x3 ->
    fill holes -> erode 4x4 -> largest 8-connected component.  PIL in -> PIL RGB 0/255 out (as the reference);
    uint8 [H, W, 3] array in -> bool [H, W] out; uint8 CUDA tensor [B, H, W, 3] in -> bit-packed int32 tensor out
    (the layout ``BatchGenerator.generate(masks=...)`` takes)."""
    if isinstance(image, torch.Tensor):
        return obj_mask_batch(image)
    a, is_pil = _to_u8_hwc(image)
    dev = _device(device)
    bits = obj_mask_batch(torch.from_numpy(a[None]).to(dev))
    m = unpack_mask_bits(bits.cpu().numpy(), a.shape[1])[0]
    if is_pil:
        from PIL import Image
        return Image.fromarray(m).convert("RGB")
    return m


# ---------------------------------------------------------------------- check_color_similarity
def check_color_similarity(patch, defect, device=None) -> float:
    """dataset_generator.py:147-159 on the device: cosine similarity of the mean colours of two images (PIL or uint8
    [H, W, 3] arrays of any sizes), float64.  The dataset path takes the same decision inside ``ssad_generate``
    (``cosine > 0.99`` -> the patch is brightened twice, datasets.py:295-299)."""
    import ctypes as C
    from . import _native as N
    a, _ = _to_u8_hwc(patch)
    b, _ = _to_u8_hwc(defect)
    dev = _device(device)
    lib = N.load()
    with torch.cuda.device(dev):
        ta, tb = torch.from_numpy(np.ascontiguousarray(a)).to(dev), torch.from_numpy(np.ascontiguousarray(b)).to(dev)
        out = torch.zeros((1,), dtype=torch.float64, device=dev)
        stream = torch.cuda.current_stream(dev).cuda_stream
        N.check(lib.ssad_color_similarity(C.c_void_p(ta.data_ptr()), a.shape[0] * a.shape[1], C.c_void_p(tb.data_ptr()),
                                          b.shape[0] * b.shape[1], C.c_void_p(out.data_ptr()), C.c_void_p(stream)),
                "ssad_color_similarity")
        return float(out.item())


# ---------------------------------------------------------------------- generate_patch / rect2poly / paste_patch
def generate_patch(image, area_ratio: Tuple[float, float] = (0.02, 0.15),
                   aspect_ratio: Tuple[Tuple[float, float], Tuple[float, float]] = ((0.3, 1), (1, 3.3)),
                   augs=None, colorized: bool = False, color_type: str = "random",
                   rng: Optional[np.random.Generator] = None):
    """dataset_generator.py:164-210: rectangle of ``area_ratio`` / ``aspect_ratio`` at a uniform position; its content
    is the crop (colorized=False), the crop's ``int(mean)`` colour ('average') or a random colour ('random').
    Host-side: this is parameter sampling plus a crop of at most a few thousand pixels (reference: 0.02 ms).
    ``augs`` keeps the reference's positional slot; the reference never uses it."""
    a, is_pil = _to_u8_hwc(image)
    rng = np.random.default_rng() if rng is None else rng
    H, W = a.shape[:2]
    area_u = rng.uniform(*area_ratio)
    asp = rng.uniform(*aspect_ratio[0]) if rng.integers(0, 2) == 0 else rng.uniform(*aspect_ratio[1])
    pw, ph = patch_size_from_draws(W, H, np.asarray([area_u]), np.asarray([asp]))
    pw, ph = int(pw[0]), int(ph[0])
    left, top = int(rng.integers(0, max(W - pw, 1) + 1)), int(rng.integers(0, max(H - ph, 1) + 1))
    crop = np.zeros((ph, pw, 3), dtype=np.uint8)                  # PIL's crop pads with 0 outside the image
    src = a[top:top + ph, left:left + pw]
    crop[:src.shape[0], :src.shape[1]] = src
    if colorized:
        if color_type == "average":
            crop[:] = (crop.reshape(-1, 3).astype(np.int64).sum(0) // (pw * ph)).astype(np.uint8)
        elif color_type == "random":
            crop[:] = rng.integers(0, 256, 3).astype(np.uint8)
        else:
            raise ValueError("color_type 'sample' is not on the reference's path (SURVEY.md section 8a)")
    return _from_u8_hwc(crop, is_pil)


def rect2poly(patch, regular: bool = False, sides: int = 8, points: Optional[Sequence[Tuple[int, int]]] = None,
              rng: Optional[np.random.Generator] = None, device=None):
    """dataset_generator.py:42-101, the branch the dataset takes (regular=False, sides=8): 4-8 vertices on the sides of
    the patch rectangle, ImageDraw.polygon(fill='white') without anti-aliasing -> mask.  ``points`` gives the vertices
    explicitly (else they are drawn as the reference draws them).  Returns a PIL 'RGBA' mask for a PIL patch (alpha and
    colour 255 inside, 0 outside, as the reference's canvas), else a uint8 [ph, pw] array of 0 / 255."""
    if regular or sides == 4:
        raise NotImplementedError("only the branch on the reference's path is built (regular=False, sides=8)")
    is_pil = hasattr(patch, "size") and not isinstance(patch, np.ndarray)
    pw, ph = (patch.size if is_pil else (np.asarray(patch).shape[1], np.asarray(patch).shape[0]))
    if points is None:
        rng = np.random.default_rng() if rng is None else rng
        points = polygon_vertices(pw, ph, *_redraw_vertices(pw, ph, rng))
    pts = np.zeros((1, 8, 2), dtype=np.int16)
    pts[0, :len(points)] = np.asarray(points, dtype=np.int16)
    dev = _device(device)
    lib = N.load()
    d_pts = torch.from_numpy(pts).to(dev)
    d_n = torch.tensor([len(points)], dtype=torch.uint8, device=dev)
    d_size = torch.tensor([[pw, ph]], dtype=torch.int16, device=dev)
    out = torch.zeros((1, ph, pw), dtype=torch.uint8, device=dev)
    stream = torch.cuda.current_stream(dev).cuda_stream
    N.check(lib.ssad_polygon_mask(1, C.c_void_p(d_pts.data_ptr()), C.c_void_p(d_n.data_ptr()), C.c_void_p(d_size.data_ptr()),
                                  C.c_void_p(out.data_ptr()), pw, ph, C.c_void_p(stream)), "ssad_polygon_mask")
    m = out[0].cpu().numpy()
    if is_pil:
        from PIL import Image
        return Image.fromarray(np.repeat(m[:, :, None], 4, axis=2), mode="RGBA")
    return m


def _redraw_vertices(pw: int, ph: int, rng: np.random.Generator):
    """The per-side draws of dataset_generator.py:65-98 for one patch -> (n_side[4], v[4][2])."""
    hh, hw = int(ph / 2), int(pw / 2)
    ns = rng.integers(1, 3, 4)
    v = np.zeros((4, 2), dtype=np.int64)
    lo_hi = [((hh + 1, ph), (1, hh), (1, ph)), ((1, hw), (hw + 1, pw), (1, pw)),
             ((1, hh), (hh + 1, ph), (1, ph)), ((hw + 1, pw), (1, hw), (1, pw))]
    for s in range(4):
        first, second, single = lo_hi[s]
        if ns[s] == 2:
            v[s, 0] = rng.integers(first[0], first[1] + 1)
            v[s, 1] = rng.integers(second[0], second[1] + 1)
        else:
            v[s, 0] = rng.integers(single[0], single[1] + 1)
    return ns, v


def paste_patch(image, patch, coords: tuple, mask=None, device=None):
    """dataset_generator.py:268-275: ``image.copy().paste(patch, coords, mask)`` on the device.  ``patch`` RGB or RGBA;
    ``mask`` None, an RGBA image (its alpha band is used, as Pillow does) or a single-band 0/255 image."""
    a, is_pil = _to_u8_hwc(image)
    p = np.asarray(patch)
    if p.ndim != 3 or p.shape[2] not in (3, 4):
        raise ValueError("patch must be RGB or RGBA")
    m = None
    if mask is not None:
        m = np.asarray(mask)
        if m.ndim == 3:
            m = m[..., 3] if m.shape[2] == 4 else m[..., 0]
        if m.shape != p.shape[:2]:
            raise ValueError("images do not match")            # Pillow's own message
        m = np.ascontiguousarray(m.astype(np.uint8) * 255 if m.dtype == bool else m.astype(np.uint8))   # mode '1': set = opaque
    elif p.shape[2] == 4:
        p = p[..., :3]                                           # paste without a mask ignores the source alpha
    dev = _device(device)
    lib = N.load()
    with torch.cuda.device(dev):
        d_img = torch.from_numpy(a.copy()).to(dev)
        d_patch = torch.from_numpy(np.ascontiguousarray(p)).to(dev)
        d_mask = torch.from_numpy(m).to(dev) if m is not None else None
        stream = torch.cuda.current_stream(dev).cuda_stream
        N.check(lib.ssad_paste(C.c_void_p(d_img.data_ptr()), a.shape[1], a.shape[0], C.c_void_p(d_patch.data_ptr()),
                               p.shape[1], p.shape[0], p.shape[2], C.c_void_p(d_mask.data_ptr()) if d_mask is not None else None,
                               int(coords[0]), int(coords[1]), C.c_void_p(stream)), "ssad_paste")
        return _from_u8_hwc(d_img.cpu().numpy(), is_pil)


# ---------------------------------------------------------------------- do_patch / do_scar
def _script_config(imsize, patch_loc: bool) -> GeneratorConfig:
    """The exerciser script's call shape (test_artificial_transformations.py:20-24, 72-76): image-sized working image,
    no pre-warp, no jitter; container factor and area ratios switch on ``patch_loc``."""
    return GeneratorConfig(subject="bottle", imsize=imsize, jitter=False, patch_localization=False, force_warp=False,
                           num_labels=4, force_line_patch_mode=bool(patch_loc),
                           force_container_factor=1 if patch_loc else None,
                           patch_area_ratio=CPP.rectangle_area_ratio_patch if patch_loc else None,
                           scar_area_ratio=CPP.scar_area_ratio_patch if patch_loc else None)


def _do_defect(label: int, img, image_for_cutting, segmentation, patch_loc: bool, params: Optional[ParamBatch],
               rng: Optional[np.random.Generator], device):
    a, is_pil = _to_u8_hwc(img)
    H, W = a.shape[:2]
    cfg = _script_config((W, H), patch_loc)
    if params is None:
        params = sample_params(cfg, 1, rng, labels=np.full(1, label, dtype=np.int64))
    elif len(params) != 1 or int(params.records["label"][0]) != label:
        raise ValueError(f"params must hold one record with label {label}")
    mask = _mask_bool(segmentation)
    dev = _device(device)
    donors = None
    if image_for_cutting is not None and image_for_cutting is not img:
        d, _ = _to_u8_hwc(image_for_cutting)
        if d.shape != a.shape:
            raise ValueError("image_for_cutting must have the size of img")
        if not np.array_equal(d, a):
            # the cut comes from another image: run it as a one-image donor bank (texture path, datasets.py:225-226)
            cfg = GeneratorConfig(**{**cfg.__dict__, "subject": "carpet", "n_donors": 1})
            params = ParamBatch(cfg, params.draws, params.records.copy())
            params.records["donor_idx"] = 0
            donors = d[None]
    if mask is not None and mask.all():
        mask = None                                              # all-white segmentation == SSAD_MASK_ALL
    gen = BatchGenerator(cfg, dev, mask=mask, donors=donors)
    out = gen.generate(torch.from_numpy(a[None]).to(dev), params, want_u8=True)
    return _from_u8_hwc(out["x_u8"][0].cpu().numpy(), is_pil)


def do_patch(img, image_for_cutting=None, segmentation=None, patch_loc: bool = False, params: Optional[ParamBatch] = None,
             rng: Optional[np.random.Generator] = None, device=None):
    """test_artificial_transformations.py:20-69: big convex 4-8-gon cut-paste (label 1) on one image."""
    return _do_defect(1, img, image_for_cutting, segmentation, patch_loc, params, rng, device)


def do_scar(img, image_for_cutting=None, segmentation=None, patch_loc: bool = False, params: Optional[ParamBatch] = None,
            rng: Optional[np.random.Generator] = None, device=None):
    """test_artificial_transformations.py:72-122: small rectangle rotated in [-45, 45] degrees, pasted 2-5 times (label 2)."""
    return _do_defect(2, img, image_for_cutting, segmentation, patch_loc, params, rng, device)


def do_swirl(img, image_for_cutting=None, segmentation=None, patch_loc: bool = False, subject=None,
             params: Optional[ParamBatch] = None, rng: Optional[np.random.Generator] = None, device=None):
    """test_artificial_transformations.py:125-164 (the dataset's label 3, datasets.py:357-388): a smoothed poly-line
    through picks of the mask, drawn black / white / silver, width 3 (1 and all 30 points with ``patch_loc``)."""
    return _do_defect(3, img, image_for_cutting, segmentation, patch_loc, params, rng, device)
