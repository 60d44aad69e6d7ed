"""GPU parity of the unit entry points of the C ABI against Pillow / the numpy model."""
import ctypes as C
import random

import numpy as np
import pytest
import torch

pytestmark = pytest.mark.gpu


def _lib():
    from ssad_b200 import _native as N
    return N, N.load()


def _p(t):
    return C.c_void_p(t.data_ptr())


def test_polygon_mask_matches_pillow():
    from PIL import Image, ImageDraw
    from test_pillow_model import _rect2poly_points
    N, lib = _lib()
    rnd = random.Random(3)
    n, MW, MH = 600, 160, 150
    xy = np.zeros((n, 8, 2), np.int16); pn = np.zeros(n, np.uint8); sz = np.zeros((n, 2), np.int16)
    refs = []
    for i in range(n):
        w, h = rnd.randint(2, MW if i % 3 else 12), rnd.randint(2, MH if i % 3 else 12)
        pts = _rect2poly_points(w, h, rnd)
        xy[i, :len(pts)] = pts; pn[i] = len(pts); sz[i] = (w, h)
        m = Image.new('RGBA', (w, h), (0, 0, 0, 0))
        ImageDraw.Draw(m).polygon(pts, fill='white')
        refs.append(np.array(m)[..., 3])
    out = torch.zeros((n, MH, MW), dtype=torch.uint8, device="cuda")
    dxy, dpn, dsz = torch.from_numpy(xy).cuda(), torch.from_numpy(pn).cuda(), torch.from_numpy(sz).cuda()
    N.check(lib.ssad_polygon_mask(n, _p(dxy), _p(dpn), _p(dsz), _p(out), MW, MH, None), "polygon")
    out = out.cpu().numpy()
    for i in range(n):
        w, h = sz[i]
        assert np.array_equal(out[i, :h, :w], refs[i]), (i, sz[i], xy[i, :pn[i]].tolist())
        assert out[i, h:].sum() == 0 and out[i, :, w:].sum() == 0


def test_affine_nearest_matches_pillow_rotate_and_random_affine():
    from PIL import Image
    from torchvision.transforms import functional as F, InterpolationMode
    from oracle import pillow_model as pm
    N, lib = _lib()
    rnd = random.Random(4); rng = np.random.default_rng(4)
    for t in range(120):
        w, h = rnd.randint(2, 60), rnd.randint(2, 60)
        angle = rnd.randint(-45, 45)
        src = rng.integers(0, 256, (h, w, 4), dtype=np.uint8); src[..., 3] = 255
        ref = np.array(Image.fromarray(src, 'RGBA').rotate(angle, expand=True))
        m, nw, nh = pm.rotate_expand_matrix(w, h, angle)
        fix = (C.c_int32 * 6)(*pm.affine_fixed_coeffs(m))
        out = torch.zeros((nh, nw, 4), dtype=torch.uint8, device="cuda")
        dsrc = torch.from_numpy(src).cuda()
        N.check(lib.ssad_affine_nearest(_p(dsrc), w, h, 4, fix, _p(out), nw, nh, None), "affine")
        assert np.array_equal(out.cpu().numpy(), ref), (w, h, angle)
    for t in range(30):
        w, h = int(rng.integers(16, 300)), int(rng.integers(16, 300))
        angle = float(np.float32(rng.uniform(-3, 3))); scale = float(np.float32(rng.uniform(1.05, 1.1)))
        src = rng.integers(0, 256, (h, w, 3), dtype=np.uint8)
        ref = np.array(F.affine(Image.fromarray(src), angle, (0, 0), scale, (0.0, 0.0),
                                interpolation=InterpolationMode.NEAREST, fill=0, center=None))
        fix = (C.c_int32 * 6)(*pm.affine_fixed_coeffs(pm.random_affine_matrix(w, h, angle, scale)))
        out = torch.zeros((h, w, 3), dtype=torch.uint8, device="cuda")
        dsrc = torch.from_numpy(src).cuda()
        N.check(lib.ssad_affine_nearest(_p(dsrc), w, h, 3, fix, _p(out), w, h, None), "affine")
        assert np.array_equal(out.cpu().numpy(), ref)


def test_color_jitter_matches_torchvision_all_orders():
    import itertools
    from PIL import Image
    from torchvision.transforms import functional as F
    N, lib = _lib()
    rng = np.random.default_rng(5)
    perms = list(itertools.permutations(range(4)))
    n, H, W = 96, 24, 40
    imgs = rng.integers(0, 256, (n, H, W, 3), dtype=np.uint8)
    imgs[::3] = (imgs[::3] // 16 + 100)
    imgs[1::7] = 255
    order = np.array([perms[i % 24] for i in range(n)], np.uint8)
    fac = np.zeros((n, 4), np.float32); fac[:, :3] = rng.uniform(0.7, 1.3, (n, 3)).astype(np.float32)
    fac[5] = 1.0
    en = np.full(n, 7, np.uint8); en[3::11] = 5; en[4::11] = 2
    out = torch.zeros((n, H, W, 3), dtype=torch.uint8, device="cuda")
    dimgs, dorder, dfac, den = (torch.from_numpy(a).cuda() for a in (imgs, order, fac, en))
    N.check(lib.ssad_color_jitter(n, H, W, _p(dimgs), _p(out), _p(dorder), _p(dfac), _p(den), None), "jitter")
    out = out.cpu().numpy()
    for i in range(n):
        x = Image.fromarray(imgs[i])
        for fn in order[i]:
            if fn == 0 and en[i] & 1: x = F.adjust_brightness(x, float(fac[i, 0]))
            elif fn == 1 and en[i] & 2: x = F.adjust_contrast(x, float(fac[i, 1]))
            elif fn == 2 and en[i] & 4: x = F.adjust_saturation(x, float(fac[i, 2]))
        assert np.array_equal(out[i], np.array(x)), i


def test_select_kth_matches_coords_map():
    from ssad_b200.generator import pack_mask_bits
    N, lib = _lib()
    rng = np.random.default_rng(6)
    for t in range(12):
        H, W = int(rng.integers(40, 300)), int(rng.integers(40, 300))
        m = rng.random((H, W)) < (0.02 if t % 3 == 0 else 0.5)
        wx, wy = int(rng.integers(0, W // 2)), int(rng.integers(0, H // 2))
        ww, wh = int(rng.integers(1, W - wx + 1)), int(rng.integers(1, H - wy + 1))
        win = m[wy:wy + wh, wx:wx + ww]
        coords = np.flip(np.column_stack(np.where(win == 1)), axis=1)     # datasets.py:264
        cnt = len(coords)
        ks = np.concatenate([rng.integers(0, max(cnt, 1), 40), [0, max(cnt - 1, 0), cnt, cnt + 5]]).astype(np.uint32)
        bits = torch.from_numpy(pack_mask_bits(m).view(np.int32)).cuda()
        out = torch.zeros((len(ks), 2), dtype=torch.int32, device="cuda")
        c = torch.zeros(1, dtype=torch.int32, device="cuda")
        dks = torch.from_numpy(ks.view(np.int32)).cuda()
        N.check(lib.ssad_select_kth(_p(bits), bits.shape[1], wx, wy, ww, wh, len(ks), _p(dks), _p(out), _p(c), None), "select")
        assert int(c.item()) == cnt
        out = out.cpu().numpy()
        for i, k in enumerate(ks):
            exp = tuple(coords[k]) if k < cnt else (-1, -1)
            assert tuple(out[i]) == exp


def test_to_tensor_matches_torchvision():
    from torchvision import transforms
    from PIL import Image
    N, lib = _lib()
    rng = np.random.default_rng(7)
    img = rng.integers(0, 256, (3, 37, 53, 3), dtype=np.uint8)
    out = torch.zeros((3, 3, 37, 53), dtype=torch.float32, device="cuda")
    dimg = torch.from_numpy(img).cuda()
    N.check(lib.ssad_to_tensor(3, 37, 53, _p(dimg), 53 * 3, 37 * 53 * 3, _p(out), None), "to_tensor")
    for i in range(3):
        assert torch.equal(out[i].cpu(), transforms.ToTensor()(Image.fromarray(img[i])))


def test_argument_validation_raises():
    from ssad_b200 import GeneratorConfig, sample_params
    from ssad_b200.generator import BatchGenerator
    cfg = GeneratorConfig(subject="bottle", imsize=(64, 64))
    gen = BatchGenerator(cfg, "cuda", mask=np.ones((64, 64), bool))
    p = sample_params(cfg, 2, np.random.default_rng(0))
    with pytest.raises(ValueError):
        gen.generate(torch.zeros((3, 64, 64, 3), dtype=torch.uint8, device="cuda"), p)
    with pytest.raises(ValueError):
        gen.generate(torch.zeros((2, 32, 64, 3), dtype=torch.uint8, device="cuda"), p)
    with pytest.raises(ValueError):
        gen.generate(torch.zeros((2, 64, 64, 3), dtype=torch.float32, device="cuda"), p)


def test_line_mask_matches_pillow():
    """ImageDraw.line(points, fill, width) for widths 1 and 3 (label 3, datasets.py:383-388): random poly-lines incl.
    points outside the canvas, repeated points, axis-parallel and diagonal segments -- the set pixels, bit-exact."""
    import ctypes as C
    from PIL import Image, ImageDraw
    from ssad_b200 import _native as N
    lib = N.load()
    rng = np.random.default_rng(17)
    W, H, n, MAXP = 96, 80, 300, 32
    xy = np.zeros((n, MAXP, 2), dtype=np.int16)
    npts = np.zeros(n, dtype=np.uint8)
    for i in range(n):
        k = int(rng.integers(2, 31))
        pts = np.stack([rng.integers(-12, W + 12, k), rng.integers(-12, H + 12, k)], axis=1)
        if i % 4 == 0:
            pts[1] = pts[0]                               # zero-length segment
        if i % 5 == 0:
            pts[:, 1] = pts[0, 1]                         # horizontal poly-line
        if i % 7 == 0 and k >= 4:
            pts[2:4, 0] = pts[2, 0]                       # a vertical segment
        xy[i, :k], npts[i] = pts, k
    for width in (1, 3):
        d_xy, d_n = torch.from_numpy(xy).cuda(), torch.from_numpy(npts).cuda()
        out = torch.zeros((n, H, W), dtype=torch.uint8, device="cuda")
        N.check(lib.ssad_line_mask(n, C.c_void_p(d_xy.data_ptr()), C.c_void_p(d_n.data_ptr()), MAXP, width,
                                   C.c_void_p(out.data_ptr()), W, H, None), "ssad_line_mask")
        got = out.cpu().numpy() > 0
        for i in range(n):
            im = Image.new("L", (W, H), 0)
            ImageDraw.Draw(im).line([tuple(map(float, p)) for p in xy[i, :npts[i]]], fill=255, width=width)
            assert np.array_equal(got[i], np.array(im) > 0), (width, i, xy[i, :npts[i]].tolist())
