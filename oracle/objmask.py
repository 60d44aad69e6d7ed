"""CPU restatement of ``obj_mask`` (reference dataset_generator.py:27-39).

TEST INFRASTRUCTURE ONLY (see ``oracle/__init__.py``).

PARITY UNPINNED: ``obj_mask`` calls ``skimage.feature.canny`` and
``skimage.morphology.label``; scikit-image is NOT installed in this image and the
reference pins no version (no requirements / lock file), so the two skimage
routines are restated here from their published algorithm (scikit-image 0.19-0.22:
``skimage/feature/_canny.py`` + ``_canny_cy.pyx``; ``skimage/measure/_label.py``)
and cannot be checked against the library.  The SciPy ``ndimage`` steps are the
installed library, called exactly as the reference calls them.
"""
from __future__ import annotations

import numpy as np
from scipy import ndimage as ndi
from scipy.ndimage import binary_closing, binary_dilation, binary_erosion


def _gaussian_unit(image_f64: np.ndarray, sigma: float) -> np.ndarray:
    """``skimage.filters.gaussian(image, sigma, mode='constant', cval=0,
    preserve_range=False)`` for an already-float image (truncate = 4.0)."""
    return ndi.gaussian_filter(image_f64, sigma, mode='constant', cval=0.0, truncate=4.0)


def canny(image: np.ndarray, sigma: float = 1.0, low_threshold=None, high_threshold=None) -> np.ndarray:
    """``skimage.feature.canny`` (2-D, no mask, no quantiles, mode='constant').

    uint8 input: thresholds are divided by the dtype max (255); the image is
    converted to float64 in [0,1] (``img_as_float``)."""
    image = np.asarray(image)
    if image.dtype == np.uint8:
        dtype_max = 255.0
        img = image.astype(np.float64) / 255.0
    else:
        dtype_max = 1.0
        img = image.astype(np.float64)
    low = 0.1 if low_threshold is None else low_threshold / dtype_max
    high = 0.2 if high_threshold is None else high_threshold / dtype_max
    # _preprocess: smoothing with bleed-over correction
    ones = np.ones(img.shape, dtype=np.float64)
    bleed_over = _gaussian_unit(ones, sigma) + np.finfo(np.float64).eps
    smoothed = _gaussian_unit(img, sigma)
    smoothed = smoothed / bleed_over
    eroded_mask = np.ones(img.shape, dtype=bool)
    eroded_mask[:1, :] = False
    eroded_mask[-1:, :] = False
    eroded_mask[:, :1] = False
    eroded_mask[:, -1:] = False
    # gradient
    jsobel = ndi.sobel(smoothed, axis=1)
    isobel = ndi.sobel(smoothed, axis=0)
    magnitude = isobel * isobel
    magnitude += jsobel * jsobel
    np.sqrt(magnitude, out=magnitude)
    low_masked = nonmaximum_suppression_bilinear(isobel, jsobel, magnitude, eroded_mask, low)
    # hysteresis
    low_mask = low_masked > 0
    strel = np.ones((3, 3), bool)
    labels, count = ndi.label(low_mask, strel)
    if count == 0:
        return low_mask
    high_mask = low_mask & (low_masked >= high)
    nonzero_sums = np.unique(labels[high_mask])
    good_label = np.zeros((count + 1,), bool)
    good_label[nonzero_sums] = True
    return good_label[labels]


def nonmaximum_suppression_bilinear(isobel, jsobel, magnitude, eroded_mask, low_threshold) -> np.ndarray:
    """``_canny_cy._nonmaximum_suppression_bilinear`` vectorised: keep ``m`` where it is
    >= both neighbours interpolated along the gradient direction."""
    H, W = magnitude.shape
    out = np.zeros_like(magnitude)
    m = magnitude
    cand = eroded_mask & (m >= low_threshold)
    is_down, is_up = isobel <= 0, isobel >= 0
    is_left, is_right = jsobel <= 0, jsobel >= 0
    cond1 = (is_up & is_right) | (is_down & is_left)
    cond2 = (is_down & is_right) | (is_up & is_left)
    ai, aj = np.abs(isobel), np.abs(jsobel)

    pad = np.pad(m, 1, mode='constant')

    def nb(dx, dy):   # magnitude[x+dx, y+dy] with x = row, y = col
        return pad[1 + dx:1 + dx + H, 1 + dy:1 + dy + W]

    with np.errstate(divide='ignore', invalid='ignore'):
        w_ji = aj / ai
        w_ij = ai / aj
    # cond1 branches
    c1a = cand & cond1 & (ai > aj)
    c1b = cand & cond1 & ~(ai > aj)
    # cond2 (elif) branches
    c2a = cand & ~cond1 & cond2 & (ai < aj)
    c2b = cand & ~cond1 & cond2 & ~(ai < aj)

    def keep(sel, w, n11, n12, n21, n22):
        w = np.where(sel, w, 0.0)
        c_plus = (n12 * w + n11 * (1.0 - w)) <= m
        c_minus = (n22 * w + n21 * (1.0 - w)) <= m
        return sel & c_plus & c_minus

    k = keep(c1a, w_ji, nb(1, 0), nb(1, 1), nb(-1, 0), nb(-1, -1))
    k |= keep(c1b, w_ij, nb(0, 1), nb(1, 1), nb(0, -1), nb(-1, -1))
    k |= keep(c2a, w_ij, nb(0, 1), nb(-1, 1), nb(0, -1), nb(1, -1))
    k |= keep(c2b, w_ji, nb(-1, 0), nb(-1, 1), nb(1, 0), nb(1, -1))
    out[k] = m[k]
    return out


def label(image: np.ndarray) -> np.ndarray:
    """``skimage.measure.label(image)``: full (8-) connectivity, background 0, labels in
    raster order of first appearance; for a 0/255 image equal-valued == foreground."""
    lab, _ = ndi.label(np.asarray(image) != 0, structure=np.ones((3, 3), bool))
    return lab


def obj_mask_array(rgb: np.ndarray) -> np.ndarray:
    """dataset_generator.py:27-39 on an RGB u8 array -> bool[H, W]."""
    from .pillow_model import rgb_to_l
    gray = rgb_to_l(rgb)                                         # image.convert('L')
    edged = canny(gray, sigma=1.5, low_threshold=5, high_threshold=15)
    structure = np.ones((3, 3), dtype=np.uint8)                  # square(3)
    edged = binary_dilation(edged, structure).astype(int)
    edged = binary_closing(edged, structure)
    edged = ndi.binary_fill_holes(edged, structure).astype(int)
    structure = np.ones((4, 4), dtype=np.uint8)                  # square(4)
    edged = binary_erosion(edged, structure).astype(int)
    edged = (edged * 255).astype(np.uint8)
    labels = label(edged)
    return labels == np.argmax(np.bincount(labels.flat, weights=edged.flat))


def obj_mask(image):
    """PIL in, PIL RGB (0/255) out, as the reference."""
    from PIL import Image
    m = obj_mask_array(np.array(image.convert('RGB')))
    return Image.fromarray(m).convert('RGB')
