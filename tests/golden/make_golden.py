"""Generate the committed golden vectors from the reference's OWN code.

Run in the build container only (needs ``/root/reference``):

    python tests/golden/make_golden.py

For every scenario of ``tests/scenarios.py`` this drives the unmodified reference
``PretextTaskDataset.__getitem__`` (datasets.py:209-394) -- and, for the function-level
cases, the test script's ``do_patch`` / ``do_scar``
(test_artificial_transformations.py:20-122) -- through ``oracle/reference_replay.py``
with the scenario's explicit draws, and stores

    x      u8 [B, h, w, 3]   image after CPP.jitter_transforms (datasets.py:391)
    y      i64 [B]           label after the coverage rule (datasets.py:258-259)
    xn     f32 [2, 3, h, w]  Normalize(ToTensor(x)) for the first two images (datasets.py:392-393)
    sha    sha256 of the scenario's synthetic inputs + packed parameter records
             (tests refuse to compare if the inputs they rebuild differ)

in ``tests/golden/<scenario>.npz``.  The inputs themselves are rebuilt from seeds by
``tests/scenarios.py`` (kept out of the fixtures to keep them small).
"""
from __future__ import annotations

import hashlib
import os
import sys
import tempfile

import numpy as np

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
sys.path.insert(0, os.path.join(ROOT, "tests"))

import scenarios as S  # noqa: E402
from oracle import reference_replay as rr  # noqa: E402

BATCH = {"image_level": 10, "image_level_256": 4, "patch_level": 10, "patch_level_32": 12, "texture": 10,
         "non_fixed": 8, "empty_mask": 6, "single_pixel_mask": 6, "flat_grey": 8,
         "lines_image": 9, "lines_patch": 12, "lines_texture": 8, "screw_patch": 12}


def scenario_sha(sc: S.Scenario) -> str:
    h = hashlib.sha256()
    h.update(np.ascontiguousarray(sc.images).tobytes())
    if sc.mask is not None:
        h.update(np.ascontiguousarray(sc.mask).tobytes())
    if sc.donors is not None:
        h.update(np.ascontiguousarray(sc.donors).tobytes())
    h.update(sc.params.records.tobytes())
    return h.hexdigest()


def reference_outputs(sc: S.Scenario):
    """x, y, xn from the reference's own __getitem__ under the scripted tape."""
    from PIL import Image
    cfg = sc.cfg
    xs, ys, xns = [], [], []
    donors = [Image.fromarray(d) for d in sc.donors] if sc.donors is not None else None
    with tempfile.TemporaryDirectory() as tmp:
        for i in range(len(sc.params)):
            png = os.path.join(tmp, f"{i}.png")
            Image.fromarray(sc.images[i]).save(png)
            wm = S.window_mask(sc, i)
            n = int(wm.sum())
            draws = sc.params.image_draws(i)
            y_eff = int(draws["label"])
            if cfg.patch_localization and 3 * n < cfg.coverage_min3:
                y_eff = 0
            if sc.mask is None:
                seg = Image.new(size=cfg.imsize, mode="RGB", color="white")
            elif sc.mask.ndim == 3:
                seg = None           # the reference computes obj_mask(original) itself
            else:
                seg = S.mask_to_pil(sc.mask)
            x, y, orig = rr.run_getitem(cfg, png, draws, donors, seg, n, effective_label=y_eff)
            assert y == y_eff
            xs.append(np.array(x))
            ys.append(y)
            if i < 2:
                xn, _, _ = rr.run_getitem(cfg, png, draws, donors, seg, n, effective_label=y_eff, with_transform=True)
                xns.append(xn.numpy())
    return np.stack(xs), np.asarray(ys, dtype=np.int64), np.stack(xns)


def script_fn_outputs(seed: int = 0):
    """Function-level entry points do_patch / do_scar on the image_level scenario inputs
    (no warp, no jitter), image-level and patch_loc=True container factors."""
    from PIL import Image
    sc = S.build("image_level", 8, seed=7)
    seg = S.mask_to_pil(sc.mask)
    n = int(sc.mask.sum())
    out = {}
    for which, lab in (("do_patch", 1), ("do_scar", 2)):
        for patch_loc in (False, True):
            cfg = S.GeneratorConfig(subject="bottle", imsize=sc.cfg.imsize, jitter=False,
                                    patch_localization=False,
                                    patch_area_ratio=S.ssad_b200.CPP.rectangle_area_ratio_patch if patch_loc else None,
                                    scar_area_ratio=S.ssad_b200.CPP.scar_area_ratio_patch if patch_loc else None)
            pb = S.sample_params(cfg, len(sc.images), np.random.default_rng(31 + lab + 2 * patch_loc),
                                 labels=np.full(len(sc.images), lab))
            box = cfg.container(1 if patch_loc else (1.75 if lab == 1 else 2))
            xs = []
            for i in range(len(sc.images)):
                img = Image.fromarray(sc.images[i])
                d = pb.image_draws(i)
                # fallback centres are drawn inside the container of *this* call
                x = rr.run_script_fn(which, img, img, seg, patch_loc, d, n, box)
                xs.append(np.array(x))
            key = f"{which}_{'patch' if patch_loc else 'image'}"
            out[key + "_x"] = np.stack(xs)
            out[key + "_rec"] = pb.records.view(np.uint8).reshape(len(pb), -1)
            for k, v in pb.draws.items():
                out[key + "_d_" + k] = v
    out["sha"] = np.frombuffer(bytes.fromhex(scenario_sha(sc)), dtype=np.uint8)
    return out


def main():
    assert rr.available(), "reference checkout not found"
    only = sys.argv[1:]                      # python make_golden.py [scenario ...]: regenerate just these
    for name in (only or S.ALL):
        sc = S.build(name, BATCH[name])
        x, y, xn = reference_outputs(sc)
        np.savez_compressed(os.path.join(HERE, f"{name}.npz"), x=x, y=y, xn=xn,
                            sha=np.frombuffer(bytes.fromhex(scenario_sha(sc)), dtype=np.uint8))
        print(name, x.shape, "labels", y.tolist())
    if not only:
        np.savez_compressed(os.path.join(HERE, "script_fns.npz"), **script_fn_outputs())
        print("script_fns done")


if __name__ == "__main__":
    main()
