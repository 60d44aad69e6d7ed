"""Pin ``oracle/pillow_model.py`` (the numpy restatement the CUDA kernels are written
from) bit-exact against the installed Pillow / torchvision -- SURVEY.md App. A."""
import random

import numpy as np
import pytest
from PIL import Image, ImageDraw, ImageEnhance

from oracle import pillow_model as pm


def _rect2poly_points(w, h, rnd):
    """vertex family of dataset_generator.py:63-98"""
    pts = []
    for side in range(4):
        n = rnd.randint(1, 2)
        if side == 0:
            pts += [(0, rnd.randint(1, h))] if n == 1 else [(0, rnd.randint(int(h / 2) + 1, h)), (0, rnd.randint(1, int(h / 2)))]
        if side == 1:
            pts += [(rnd.randint(1, w), 0)] if n == 1 else [(rnd.randint(1, int(w / 2)), 0), (rnd.randint(int(w / 2) + 1, w), 0)]
        if side == 2:
            pts += [(w, rnd.randint(1, h))] if n == 1 else [(w, rnd.randint(1, int(h / 2))), (w, rnd.randint(int(h / 2) + 1, h))]
        if side == 3:
            pts += [(rnd.randint(1, w), h)] if n == 1 else [(rnd.randint(int(w / 2) + 1, w), h), (rnd.randint(1, int(w / 2)), h)]
    return pts


@pytest.mark.parametrize("maxdim,count", [(12, 600), (130, 500), (500, 60)])
def test_polygon_fill_matches_pillow(maxdim, count):
    rnd = random.Random(maxdim)
    for _ in range(count):
        w, h = rnd.randint(2, maxdim), rnd.randint(2, maxdim)
        pts = _rect2poly_points(w, h, rnd)
        m = Image.new('RGBA', (w, h), (0, 0, 0, 0))
        ImageDraw.Draw(m).polygon(pts, fill='white')
        ref = np.array(m)[..., 3]
        assert set(np.unique(ref)) <= {0, 255}
        assert np.array_equal(ref > 0, pm.polygon_mask(pts, w, h)), (w, h, pts)


def test_rotate_expand_matches_pillow():
    rnd = random.Random(5)
    rng = np.random.default_rng(5)
    for _ in range(300):
        w, h = rnd.randint(2, 60), rnd.randint(2, 60)
        angle = rnd.randint(-45, 45)
        src = rng.integers(0, 256, (h, w, 4), dtype=np.uint8)
        src[..., 3] = 255
        ref = np.array(Image.fromarray(src, 'RGBA').rotate(angle, expand=True))
        m, nw, nh = pm.rotate_expand_matrix(w, h, angle)
        assert ref.shape[:2] == (nh, nw)
        got = pm.affine_nearest_fixed(src, pm.affine_fixed_coeffs(m), nw, nh)
        assert np.array_equal(ref, got), (w, h, angle)


def test_random_affine_matches_torchvision():
    from torchvision.transforms import functional as F, InterpolationMode
    rng = np.random.default_rng(6)
    for _ in range(40):
        w, h = int(rng.integers(16, 200)), int(rng.integers(16, 200))
        angle = float(np.float32(rng.uniform(-3, 3)))
        scale = float(np.float32(rng.uniform(1.05, 1.1)))
        src = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        ref = np.array(F.affine(Image.fromarray(src), angle, (0, 0), scale, (0.0, 0.0),
                                interpolation=InterpolationMode.NEAREST, fill=0, center=None))
        m = pm.random_affine_matrix(w, h, angle, scale)
        got = pm.affine_nearest_fixed(src, pm.affine_fixed_coeffs(m), w, h)
        assert np.array_equal(ref, got)


def test_blend_and_enhancers_match_pillow():
    rng = np.random.default_rng(7)
    for _ in range(120):
        img = rng.integers(0, 256, (17, 23, 3), dtype=np.uint8)
        pil = Image.fromarray(img)
        f = float(np.float32(rng.uniform(0.7, 1.2)))
        assert np.array_equal(np.array(ImageEnhance.Brightness(pil).enhance(f)), pm.brightness(img, f))
        assert np.array_equal(np.array(ImageEnhance.Contrast(pil).enhance(f)), pm.contrast(img, f))
        assert np.array_equal(np.array(ImageEnhance.Color(pil).enhance(f)), pm.saturation(img, f))
        assert np.array_equal(np.array(pil.convert('L')), pm.rgb_to_l(img))
    # float64 factors as the reference passes to ImageEnhance.Brightness (datasets.py:296-299)
    for _ in range(60):
        img = rng.integers(0, 256, (9, 11, 3), dtype=np.uint8)
        f = float(rng.uniform(0.75, 1.15))
        assert np.array_equal(np.array(ImageEnhance.Brightness(Image.fromarray(img)).enhance(f)),
                              pm.brightness(img, float(np.float32(f))))


def test_color_jitter_chain_matches_torchvision():
    from torchvision.transforms import functional as F
    import itertools
    rng = np.random.default_rng(8)
    perms = list(itertools.permutations(range(4)))
    for t in range(96):
        img = rng.integers(0, 256, (20, 31, 3), dtype=np.uint8)
        if t % 3 == 0:   # smooth image: exercises near-mean contrast values
            img = (img // 16 + 100).astype(np.uint8)
        order = perms[t % len(perms)]
        fac = [float(np.float32(rng.uniform(0.9, 1.1))) for _ in range(3)]
        x = Image.fromarray(img)
        for fn in order:
            if fn == 0:
                x = F.adjust_brightness(x, fac[0])
            elif fn == 1:
                x = F.adjust_contrast(x, fac[1])
            elif fn == 2:
                x = F.adjust_saturation(x, fac[2])
        assert np.array_equal(np.array(x), pm.color_jitter(img, order, fac))


def test_paste_clipping_matches_pillow():
    rng = np.random.default_rng(9)
    for _ in range(100):
        H, W = 40, 50
        dst = rng.integers(0, 256, (H, W, 3), dtype=np.uint8)
        ph, pw = int(rng.integers(2, 60)), int(rng.integers(2, 60))
        src = rng.integers(0, 256, (ph, pw, 4), dtype=np.uint8)
        src[..., 3] = np.where(rng.random((ph, pw)) < 0.5, 0, 255)
        l, t = int(rng.integers(-30, 60)), int(rng.integers(-30, 50))
        d = Image.fromarray(dst).copy()
        s = Image.fromarray(src, 'RGBA')
        d.paste(s, (l, t), mask=s)
        assert np.array_equal(np.array(d), pm.paste_select(dst, src, l, t, src[..., 3]))


def test_convert_1_is_identity_on_binary_masks():
    m = np.random.default_rng(10).random((33, 47)) < 0.4
    rgb = Image.fromarray(m).convert('RGB')
    assert np.array_equal(np.array(rgb.convert('1')), m)


def test_container_and_clamp_match_reference_formulas():
    assert pm.container_box(256, 1.75) == (54, 54, 201, 201)
    assert pm.container_box(256, 2) == (64, 64, 192, 192)
    assert pm.container_box(256, 1) == (0, 0, 256, 256)
    # patch wider than the container: starts at L and overhangs (dataset_generator.py:132-143)
    assert pm.clamp_to_container((0, 0, 32, 32), (41, 10), (20, 20)) == (0, 15)


def test_cosine_of_means_matches_sklearn():
    from sklearn.metrics.pairwise import cosine_similarity
    rng = np.random.default_rng(11)
    for _ in range(200):
        a = rng.integers(0, 256, (12, 9, 3), dtype=np.uint8)
        b = rng.integers(0, 256, (5, 7, 3), dtype=np.uint8)
        ca, cb = a.mean(axis=(0, 1)) / 255, b.mean(axis=(0, 1)) / 255
        ref = float(cosine_similarity(ca[None], cb[None]).squeeze())
        got = pm.cosine_of_means(a.reshape(-1, 3).sum(0), 108, b.reshape(-1, 3).sum(0), 35)
        assert abs(ref - got) <= 1e-15
        assert (ref > 0.99) == (got > 0.99)
