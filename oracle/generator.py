"""Explicit-parameter CPU restatement of the reference's per-image generator.

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

Follows, statement by statement,
  * ``PretextTaskDataset.__getitem__``  /root/reference/src/self_supervised/datasets.py:209-394
  * the generator primitives             /root/reference/src/self_supervised/dataset_generator.py:15-275
  * ``do_patch`` / ``do_scar``           /root/reference/src/test_artificial_transformations.py:20-122
but every random draw the reference makes inside those functions is taken from an
explicit per-image record ``d`` (a dict; field list in ``DRAW_FIELDS``) instead of
``random`` / ``np.random`` / ``torch``.  All pixel arithmetic is done by the same
installed Pillow / torchvision routines the reference calls, through the same
calls, so that "the reference applied to the same parameters" is exactly this.

``tests/test_golden.py`` pins this file against outputs of the reference's OWN
code run under a scripted RNG (``oracle/reference_replay.py`` ->
``tests/golden/*.npz``).
"""
from __future__ import annotations

from typing import Dict, List, Optional, Sequence, Tuple

import numpy as np
from PIL import Image, ImageDraw, ImageEnhance

DRAW_FIELDS = (
    "label",                       # datasets.py:215
    "warp_angle", "warp_scale",    # RandomAffine.get_params, datasets.py:221-222
    "donor_idx",                   # random.choice(images_for_cut), datasets.py:226 (-1: original)
    "crop_left", "crop_top",       # datasets.py:249-250
    "donor_i", "donor_j",          # RandomCrop.get_params (top i, left j), datasets.py:253
    "centre_r",                    # u32[6]: index = (r * N) >> 32 for randint(0, N-1), dataset_generator.py:264
    "variant",                     # np.random.choice([0,1,2]), datasets.py:270,311
    "area_u", "asp0", "asp1", "asp_pick",   # dataset_generator.py:173-174
    "rect_l", "rect_t",            # dataset_generator.py:190
    "rnd_rgb",                     # dataset_generator.py:195-199
    "low", "high", "pick1", "pick2",        # datasets.py:296-299
    "fallback_c",                  # i[6][2] dataset_generator.py:113-114
    "poly_n_side", "poly_v",       # i[4], i[4][2]  dataset_generator.py:65-98
    "scar_k", "scar_angle",        # datasets.py:341-342
    "jit_order", "jit_b", "jit_c", "jit_s",  # ColorJitter.get_params, datasets.py:391
)

NORM_MEAN = (0.485, 0.456, 0.406)   # datasets.py:430-433
NORM_STD = (0.229, 0.224, 0.225)


# ---------------------------------------------------------------- primitives
class Container:
    """dataset_generator.py:15-24"""

    def __init__(self, imsize: tuple, scaling_factor: float) -> None:
        self.center = int(imsize[0] / 2)
        self.dim = int(imsize[0] / scaling_factor)
        self.left = int(self.center - (self.center / scaling_factor))
        self.top = int(self.center - (self.center / scaling_factor))
        self.right = int(self.center + (self.center / scaling_factor))
        self.bottom = int(self.center + (self.center / scaling_factor))
        self.width = self.right - self.left
        self.height = self.bottom - self.top


def patch_size_from_draws(image_size: Tuple[int, int], area_ratio_u: float,
                          asp0: float, asp1: float, asp_pick: int) -> Tuple[int, int]:
    """dataset_generator.py:172-180 -- ``area_ratio_u`` is the value
    ``random.uniform(area_ratio[0], area_ratio[1])`` returned."""
    img_area = image_size[0] * image_size[1]
    patch_area = area_ratio_u * img_area
    patch_aspect = [asp0, asp1][asp_pick]
    patch_w = int(np.sqrt(patch_area * patch_aspect))
    patch_h = int(np.sqrt(patch_area / patch_aspect))
    if patch_w < 2:
        patch_w = 2
    if patch_h < 2:
        patch_h = 2
    return patch_w, patch_h


def generate_patch(image: Image.Image, d: Dict, variant: int) -> Image.Image:
    """dataset_generator.py:164-210 with the draws taken from ``d``.
    variant 0 = cut, 1 = colorized 'average', 2 = colorized 'random'
    (datasets.py:270-293)."""
    patch_w, patch_h = patch_size_from_draws(image.size, float(d["area_u"]), float(d["asp0"]),
                                             float(d["asp1"]), int(d["asp_pick"]))
    org_w, org_h = image.size
    w = org_w - patch_w
    if w < 1:
        w = 1
    h = org_h - patch_h
    if h < 1:
        h = 1
    patch_left, patch_top = int(d["rect_l"]), int(d["rect_t"])
    assert 0 <= patch_left <= w and 0 <= patch_top <= h, "rect origin outside the reference's randint range"
    patch_right, patch_bottom = patch_left + patch_w, patch_top + patch_h
    if variant == 2:
        rgb = tuple(int(v) for v in d["rnd_rgb"])
        return Image.new('RGB', (patch_w, patch_h), color=rgb)
    if variant == 1:
        patch = image.crop((patch_left, patch_top, patch_right, patch_bottom))
        imarray = np.array(patch)
        color = imarray.mean(axis=(0, 1))
        rgb = (int(color[0]), int(color[1]), int(color[2]))
        return Image.new('RGB', (patch_w, patch_h), color=rgb)
    return image.crop((patch_left, patch_top, patch_right, patch_bottom))


def check_color_similarity(patch: Image.Image, defect: Image.Image) -> float:
    """dataset_generator.py:147-159"""
    from sklearn.metrics.pairwise import cosine_similarity
    imarray = np.array(patch)
    color = imarray.mean(axis=(0, 1))
    rgb1 = (float(color[0] / 255), float(color[1] / 255), float(color[2] / 255))
    imarray = np.array(defect)
    color = imarray.mean(axis=(0, 1))
    rgb2 = (float(color[0] / 255), float(color[1] / 255), float(color[2] / 255))
    v1 = np.array(rgb1)[None, :]
    v2 = np.array(rgb2)[None, :]
    return cosine_similarity(v1, v2).squeeze()


def polygon_points(width: int, height: int, n_side: Sequence[int], v: Sequence[Sequence[int]]) -> List[Tuple[int, int]]:
    """The vertex list ``rect2poly(patch, regular=False, sides=8)`` builds
    (dataset_generator.py:63-98) from per-side draws: ``n_side[s]`` in {1,2} and the
    one or two coordinates ``v[s]`` drawn for that side, in draw order."""
    pts: List[Tuple[int, int]] = []
    for side in range(4):
        n = int(n_side[side])
        a, b = int(v[side][0]), int(v[side][1])
        if side == 0:
            pts += [(0, a)] if n == 1 else [(0, a), (0, b)]
        elif side == 1:
            pts += [(a, 0)] if n == 1 else [(a, 0), (b, 0)]
        elif side == 2:
            pts += [(width, a)] if n == 1 else [(width, a), (width, b)]
        else:
            pts += [(a, height)] if n == 1 else [(a, height), (b, height)]
    return pts


def rect2poly(patch: Image.Image, points) -> Image.Image:
    """dataset_generator.py:42-101 (the ``else``/``sides != 4`` branch)."""
    mask = Image.new('RGBA', (patch.size), color=(0, 0, 0, 0))
    draw = ImageDraw.Draw(mask)
    draw.polygon(points, fill='white')
    return mask


def check_valid_coordinates_by_container(imsize, patchsize, current_coords, container_scaling_factor,
                                         fallback) -> Tuple[int, int]:
    """dataset_generator.py:104-144; ``fallback`` replaces the two randint draws."""
    defect_width, defect_height = patchsize
    container = Container(imsize, scaling_factor=container_scaling_factor)
    if current_coords is None:
        center_x, center_y = int(fallback[0]), int(fallback[1])
        assert container.left <= center_x <= container.right and container.top <= center_y <= container.bottom
    else:
        center_x = current_coords[0]
        center_y = current_coords[1]
    paste_left = center_x - int(patchsize[0] / 2)
    paste_top = center_y - int(patchsize[1] / 2)
    paste_right = center_x + int(patchsize[0] / 2)
    paste_bottom = center_y + int(patchsize[1] / 2)
    if paste_right > container.right:
        paste_left = container.right - defect_width
    if paste_bottom > container.bottom:
        paste_top = container.bottom - defect_height
    if paste_left < container.left:
        paste_left = container.left
    if paste_top < container.top:
        paste_top = container.top
    return (int(paste_left), int(paste_top))


def get_coordinate(xy_coords: np.ndarray, r: int):
    """dataset_generator.py:259-265 with ``randint(0, N-1) := (r * N) >> 32``."""
    n = len(xy_coords)
    if n == 0:
        return None
    elif n < 2:
        return xy_coords[0]
    idx = (int(r) * n) >> 32
    return xy_coords[idx]


def paste_patch(image: Image.Image, patch: Image.Image, coords, mask: Image.Image = None) -> Image.Image:
    """dataset_generator.py:268-275"""
    aug_image = image.copy()
    aug_image.paste(patch, (int(coords[0]), int(coords[1])), mask=mask)
    return aug_image


def coords_map_of(segmentation: Image.Image) -> np.ndarray:
    """datasets.py:263-264"""
    seg = np.array(segmentation.convert('1'))
    return np.flip(np.column_stack(np.where(seg == 1)), axis=1)


def _similarity_brightness(x: Image.Image, patch: Image.Image, d: Dict, info: Dict) -> Image.Image:
    """datasets.py:295-299 / 335-339"""
    sim = check_color_similarity(x, patch)
    info["cos"] = float(sim)
    info["sim"] = bool(sim > 0.99)
    if sim > 0.99:
        low, high = float(d["low"]), float(d["high"])
        patch = ImageEnhance.Brightness(patch).enhance([low, high][int(d["pick1"])])
        patch = ImageEnhance.Brightness(patch).enhance([low, high][int(d["pick2"])])
    return patch


# ---------------------------------------------------------------- defects
def apply_polygon(x: Image.Image, image_for_cutting: Image.Image, coords_map: np.ndarray, d: Dict,
                  area_ratio, factor: float, info: Dict) -> Image.Image:
    """label 1 -- datasets.py:267-308 (== do_patch, test_artificial_transformations.py:20-69)."""
    coords = get_coordinate(coords_map, d["centre_r"][0])
    patch = generate_patch(image_for_cutting, d, int(d["variant"]))
    patch = _similarity_brightness(x, patch, d, info)
    coords = check_valid_coordinates_by_container(x.size, patch.size, coords, factor, d["fallback_c"][0])
    info["paste"] = [coords]
    pts = polygon_points(patch.size[0], patch.size[1], d["poly_n_side"], d["poly_v"])
    mask = rect2poly(patch, pts)
    return paste_patch(x, patch, coords, mask)


def apply_scar(x: Image.Image, image_for_cutting: Image.Image, coords_map: np.ndarray, d: Dict,
               area_ratio, factor: float, info: Dict) -> Image.Image:
    """label 2 -- datasets.py:310-355 (== do_scar, test_artificial_transformations.py:72-122)."""
    scar = generate_patch(image_for_cutting, d, int(d["variant"]))
    scar = _similarity_brightness(x, scar, d, info)
    scar = scar.convert('RGBA')
    k = int(d["scar_k"])
    angle = int(d["scar_angle"])
    s = scar.rotate(angle, expand=True)
    info["paste"] = []
    info["scar_size"] = s.size
    for i in range(k):
        coords = get_coordinate(coords_map, d["centre_r"][i])
        coords = check_valid_coordinates_by_container(x.size, s.size, coords, factor, d["fallback_c"][i])
        info["paste"].append(coords)
        x = paste_patch(x, s, coords, s)
    return x


def apply_line(x: Image.Image, coords_map: np.ndarray, d: Dict, patch_localization: bool, info: Dict) -> Image.Image:
    """Label 3 "long line", datasets.py:357-388 statement by statement with the draws made explicit
    (``random.randint(a, b)`` = a + ((r * (b - a + 1)) >> 32) with the 32-bit draw r, as for the centre picks).
    An empty mask makes the reference raise IndexError (``coords_map[0]``); here the image stays as it is."""
    from PIL import ImageDraw
    from scipy.signal import savgol_filter
    if len(coords_map) == 0:
        info["line_points"] = []
        return x
    draw = ImageDraw.Draw(x)
    side = ['left', 'top'][int(d["line_side"])]
    n = 60 if not patch_localization else 30
    points = []
    c = 0
    for i in range(n):
        offset = i / n
        hi = int(len(coords_map) * offset)
        index = c + ((int(d["line_r"][i]) * (hi - c + 1)) >> 32)
        p = coords_map[index]
        points.append(tuple(p))
        c = index
    rgb = ['black', 'white', 'silver'][int(d["line_rgb"])]
    if side == 'left':
        points.sort(key=lambda tup: tup[0])
    points = savgol_filter(points, 10, 2, axis=0)
    if not patch_localization:
        p_splits = np.array_split(points, 10)
        points = p_splits[int(d["line_k"])]
    points = [tuple(p) for p in points]
    info["line_points"] = points
    draw.line(points, fill=rgb, width=1 if patch_localization else 3)
    return x


def color_jitter(x: Image.Image, d: Dict) -> Image.Image:
    """``CPP.jitter_transforms(x)`` -- datasets.py:44-47,391; torchvision
    ColorJitter.forward with explicit fn_idx and factors (hue is None)."""
    from torchvision.transforms import functional as F
    for fn_id in d["jit_order"]:
        fn_id = int(fn_id)
        if fn_id == 0:
            x = F.adjust_brightness(x, float(d["jit_b"]))
        elif fn_id == 1:
            x = F.adjust_contrast(x, float(d["jit_c"]))
        elif fn_id == 2:
            x = F.adjust_saturation(x, float(d["jit_s"]))
    return x


# ---------------------------------------------------------------- the item
def generate_item(original: Image.Image, d: Dict, cfg, donors: Optional[List[Image.Image]],
                  fixed_segmentation: Optional[Image.Image], mask_fn=None):
    """``PretextTaskDataset.__getitem__`` (datasets.py:209-394) from the decoded,
    resized RGB ``original`` onward.  Returns ``(x_pil, y, info)`` where ``x_pil`` is
    the image right after ``CPP.jitter_transforms`` (datasets.py:391).

    cfg attributes used: patch_localization, patch_size, subject, is_texture,
    is_non_fixed, patch_area_ratio, scar_area_ratio, jitter.
    """
    from torchvision.transforms import functional as F
    from torchvision.transforms import InterpolationMode
    info: Dict = {}
    y = int(d["label"])
    x = original.copy()
    if not cfg.patch_localization and not cfg.is_non_fixed:
        # transforms.RandomAffine(3, scale=(1.05,1.1)).forward with scripted get_params
        x = F.affine(x, float(d["warp_angle"]), (0, 0), float(d["warp_scale"]), (0.0, 0.0),
                     interpolation=InterpolationMode.NEAREST, fill=0, center=None)
    if cfg.is_texture:
        image_for_cutting = donors[int(d["donor_idx"])]
    else:
        image_for_cutting = original
    if cfg.is_non_fixed:
        segmentation = mask_fn(original)
    else:
        segmentation = fixed_segmentation
    factor_patch = 1.75
    factor_scar = 2
    if cfg.patch_localization:
        ps = cfg.patch_size
        if cfg.subject == 'capsule':
            x = x.crop((0, 50, 255, 200))
            segmentation = segmentation.crop((0, 50, 255, 200))
        if cfg.subject == 'screw':
            x = x.crop((25, 25, 230, 230))
            segmentation = segmentation.crop((25, 25, 230, 230))
        left, top = int(d["crop_left"]), int(d["crop_top"])
        assert 0 <= left <= x.size[0] - ps and 0 <= top <= x.size[1] - ps
        x = x.crop((left, top, left + ps, top + ps))
        segmentation = segmentation.crop((left, top, left + ps, top + ps))
        # transforms.RandomCrop(ps)(image_for_cutting) == F.crop(img, i, j, ps, ps)
        i, j = int(d["donor_i"]), int(d["donor_j"])
        image_for_cutting = F.crop(image_for_cutting, i, j, ps, ps)
        factor_patch = 1
        factor_scar = 1
        import torch
        from torchvision import transforms
        if torch.sum(transforms.ToTensor()(segmentation)) < int((ps * ps) / 2):
            y = 0
    if y > 0:
        coords_map = coords_map_of(segmentation)
        info["n_coords"] = len(coords_map)
        if y == 1:
            x = apply_polygon(x, image_for_cutting, coords_map, d, cfg.patch_area_ratio, factor_patch, info)
        elif y == 2:
            x = apply_scar(x, image_for_cutting, coords_map, d, cfg.scar_area_ratio, factor_scar, info)
        elif y == 3:
            x = apply_line(x, coords_map, d, cfg.patch_localization, info)
    if getattr(cfg, "jitter", True):
        x = color_jitter(x, d)
    return x, y, info


def to_tensor_normalize(x: Image.Image):
    """datasets.py:392-393 with the datamodule's transform (datasets.py:430-433)."""
    from torchvision import transforms
    t = transforms.Compose([transforms.ToTensor(), transforms.Normalize(NORM_MEAN, NORM_STD)])
    return t(x)


def to_tensor(original: Image.Image):
    """datasets.py:394"""
    from torchvision import transforms
    return transforms.ToTensor()(original)


# ------------------------------------------------- script-level entry points
def do_patch(img: Image.Image, image_for_cutting: Image.Image, segmentation: Image.Image, patch_loc: bool,
             d: Dict, info: Optional[Dict] = None) -> Image.Image:
    """test_artificial_transformations.py:20-69 (the unused ``k = randint(1,2)`` draw
    at :23 has no effect on the output)."""
    info = {} if info is None else info
    factor = 1 if patch_loc else 1.75
    x = img.copy()
    coordinates = coords_map_of(segmentation)
    return apply_polygon(x, image_for_cutting, coordinates, d, None, factor, info)


def do_scar(img: Image.Image, image_for_cutting: Image.Image, segmentation: Image.Image, patch_loc: bool,
            d: Dict, info: Optional[Dict] = None) -> Image.Image:
    """test_artificial_transformations.py:72-122"""
    info = {} if info is None else info
    factor = 1 if patch_loc else 2
    x = img.copy()
    coordinates = coords_map_of(segmentation)
    return apply_scar(x, image_for_cutting, coordinates, d, None, factor, info)


def do_swirl(img: Image.Image, image_for_cutting: Image.Image, segmentation: Image.Image, patch_loc: bool,
             d: Dict, info: Optional[Dict] = None) -> Image.Image:
    """test_artificial_transformations.py:125-164 (same body as the dataset's label 3)."""
    info = {} if info is None else info
    x = img.copy()
    coordinates = coords_map_of(segmentation)
    return apply_line(x, coordinates, d, patch_loc, info)
