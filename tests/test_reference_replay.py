"""Container-only checks against the reference checkout (skipped where
``/root/reference`` is absent, e.g. on the GPU box):

* the scripted tape is transparent: the reference run UNSCRIPTED under seeded global
  RNGs, with its draws recorded, is reproduced byte-for-byte by replaying the record;
* the explicit-parameter oracle equals the replayed reference on fresh random draws.
"""
import os
import random
import tempfile

import numpy as np
import pytest

import scenarios as S
from oracle import reference_replay as rr

pytestmark = pytest.mark.skipif(not rr.available(), reason="reference checkout not present")


def _blank_draws():
    from ssad_b200.params import empty_draws
    d = empty_draws(1)
    return {k: (v[0].copy() if v[0].ndim else v[0].item()) for k, v in d.items()}


@pytest.mark.parametrize("name", ["image_level", "patch_level", "texture", "empty_mask", "flat_grey", "lines_image", "lines_patch"])
def test_record_then_replay_is_transparent(name):
    import torch
    from PIL import Image
    sc = S.build(name, 6)
    cfg = sc.cfg
    donors = [Image.fromarray(d) for d in sc.donors] if sc.donors is not None else None
    seg = Image.new(size=cfg.imsize, mode="RGB", color="white") if sc.mask is None else S.mask_to_pil(sc.mask)
    with tempfile.TemporaryDirectory() as tmp:
        for i in range(len(sc.images)):
            png = os.path.join(tmp, f"{i}.png")
            Image.fromarray(sc.images[i]).save(png)
            # unscripted run, seeded like tools.py:232-238, with the draws recorded
            for attempt in range(50):
                random.seed(100 * i + attempt)
                np.random.seed(100 * i + attempt)
                torch.manual_seed(100 * i + attempt)
                st = (random.getstate(), np.random.get_state(), torch.get_rng_state())
                lab0 = random.randint(0, 3)
                if name.startswith("lines"):        # HEAD's 4-label stream: take the seeds that draw label 3 ("long line")
                    if lab0 == 3 or attempt >= 8:
                        break
                elif lab0 != 3:                     # 3-class contract: skip seeds drawing label 3
                    break
            random.setstate(st[0]); np.random.set_state(st[1]); torch.set_rng_state(st[2])
            d = _blank_draws()
            # n_coords: the window depends on crop draws not known yet -> find by a dry record pass
            if cfg.patch_localization:
                probe = _blank_draws()
                random.setstate(st[0])
                random.randint(0, 3)
                pc = cfg.precrop
                left = random.randint(0, pc[2] - pc[0] - cfg.patch_size)
                top = random.randint(0, pc[3] - pc[1] - cfg.patch_size)
                wm = sc.mask[pc[1] + top:pc[1] + top + cfg.patch_size, pc[0] + left:pc[0] + left + cfg.patch_size]
                random.setstate(st[0])
            else:
                wm = np.ones(cfg.imsize[::-1], bool) if sc.mask is None else sc.mask
            n = int(wm.sum())
            random.setstate(st[0]); np.random.set_state(st[1]); torch.set_rng_state(st[2])
            lab = random.randint(0, 3)
            random.setstate(st[0])
            y_eff = 0 if (cfg.patch_localization and 3 * n < cfg.coverage_min3) else lab
            x1, y1, _ = rr.run_getitem(cfg, png, d, donors, seg, n, effective_label=y_eff, record=True)
            x2, y2, _ = rr.run_getitem(cfg, png, d, donors, seg, n, effective_label=y_eff, record=False)
            assert y1 == y2
            assert np.array_equal(np.array(x1), np.array(x2))
            # and the explicit-parameter oracle on the recorded draws
            from oracle import generator as og
            x3, y3, _ = og.generate_item(Image.fromarray(sc.images[i]), d, cfg, donors, seg)
            assert y3 == y1 and np.array_equal(np.array(x3), np.array(x1))


def test_reference_obj_mask_body_equals_restatement():
    from PIL import Image
    from oracle import objmask
    from ssad_b200 import synth
    for im in synth.screw_images(3, 96, 96, 3):
        a = np.array(rr.reference_obj_mask(Image.fromarray(im)))[..., 0] > 0
        assert np.array_equal(a, objmask.obj_mask_array(im))
