"""Run the reference's OWN generator code under a scripted (or recorded) RNG.

TEST INFRASTRUCTURE ONLY, and usable only where the reference checkout exists
(``/root/reference`` in the build container; it does not exist on the GPU box).
``tests/golden/make_golden.py`` uses this to produce the committed golden vectors;
``tests/test_reference_replay.py`` (skipped when the checkout is absent) re-checks
the explicit-parameter oracle against it.

How: the reference draws randomness *inside* its functions from module-level
``random``, ``np.random`` and torchvision ``get_params``.  We import the reference
unmodified (``/root/reference/src`` on ``sys.path``; stub modules for the absent
``skimage`` / ``pytorch_lightning`` / ``matplotlib``), then replace, *as seen from
the reference's modules only*:

* ``dataset_generator.random`` / ``datasets.random`` / the test script's ``random``
* ``np.random`` as seen by ``datasets`` and the test script (a proxy module)
* ``transforms.RandomAffine.get_params``, ``RandomCrop.get_params``,
  ``ColorJitter.get_params``

by a tape that serves each draw, in the exact call order of SURVEY.md App. A.9,
from an explicit draws record -- or, in *record* mode, performs the real draw and
writes it into the record.  record -> replay reproducing the unscripted output
byte for byte proves the tape transparent.  The pixel work is done by the
reference's code + the installed Pillow/torchvision in every mode.
"""
from __future__ import annotations

import contextlib
import os
import random as _py_random
import sys
import types
from typing import Callable, Dict, Optional

import numpy as np

REFERENCE_SRC = "/root/reference/src"


def available() -> bool:
    return os.path.isdir(os.path.join(REFERENCE_SRC, "self_supervised"))


# ------------------------------------------------------------------ import
_ref_modules = None


def _install_stubs():
    from . import objmask as om

    def mod(name, **attrs):
        m = types.ModuleType(name)
        m.__dict__.update(attrs)
        sys.modules[name] = m
        return m

    if "skimage" not in sys.modules:
        feature = mod("skimage.feature", canny=om.canny)
        morphology = mod("skimage.morphology", square=lambda n: np.ones((n, n), dtype=np.uint8), label=om.label)
        segmentation = mod("skimage.segmentation", slic=None)
        color = mod("skimage.color")
        transform = mod("skimage.transform", swirl=None)
        measure = mod("skimage.measure", label=om.label)
        mod("skimage", feature=feature, morphology=morphology, segmentation=segmentation, color=color,
            transform=transform, measure=measure)
    if "pytorch_lightning" not in sys.modules:
        class LightningDataModule:
            def __init__(self, *a, **k):
                pass

            def save_hyperparameters(self, *a, **k):
                pass

        class LightningModule:
            pass

        class Callback:
            pass
        mod("pytorch_lightning", LightningDataModule=LightningDataModule, LightningModule=LightningModule,
            Callback=Callback)
    if "matplotlib" not in sys.modules:
        plt = mod("matplotlib.pyplot")
        mod("matplotlib", pyplot=plt)


def load_reference():
    """Import the reference modules (unmodified sources, from where they lie)."""
    global _ref_modules
    if _ref_modules is not None:
        return _ref_modules
    if not available():
        raise RuntimeError("reference checkout not present")
    _install_stubs()
    sys.path.insert(0, REFERENCE_SRC)
    try:
        import self_supervised.dataset_generator as dg
        import self_supervised.datasets as ds
        import test_artificial_transformations as tat
    finally:
        sys.path.remove(REFERENCE_SRC)
    _ref_modules = types.SimpleNamespace(dg=dg, ds=ds, tat=tat)
    return _ref_modules


# ------------------------------------------------------------------ the tape
class _NpProxy:
    """stands in for the ``np`` module inside a reference module: everything is numpy's
    except ``.random``."""

    def __init__(self, rnd):
        self.random = rnd

    def __getattr__(self, k):
        return getattr(np, k)


class Tape:
    """Serves (replay) or captures (record) the draws of one reference call."""

    def __init__(self, program, draws: Dict, record: bool):
        self.draws = draws
        self.record = record
        self._gen = program(self)
        next(self._gen)

    # every scripted call funnels through here
    def _call(self, name: str, args, real: Callable):
        return self._gen.send((name, args, real))

    # -- value helpers used by programs
    def val(self, req, field, idx=None, want=None, check=None, to_draw=None, from_draw=None):
        """Answer request ``req`` from ``draws[field][idx]`` (replay) or perform the real
        draw and store it (record)."""
        name, args, real = req
        if want is not None:
            assert name == want, f"draw order: expected {want} for {field}, got {name}{args}"
        if check is not None:
            assert check(args), f"draw order: unexpected arguments {name}{args} for {field}"
        if self.record:
            v = real(*args)
            sv = to_draw(v, args) if to_draw else v
            if idx is None:
                self.draws[field] = sv
            else:
                self.draws[field][idx] = sv
            return v
        sv = self.draws[field] if idx is None else self.draws[field][idx]
        return from_draw(sv, args) if from_draw else sv


class _RandomFacade:
    """Replaces the ``random`` module object inside reference modules."""

    def __init__(self, tape: Tape, real=_py_random):
        self._t, self._r = tape, real

    def randint(self, a, b):
        return self._t._call("randint", (a, b), self._r.randint)

    def uniform(self, a, b):
        return self._t._call("uniform", (a, b), self._r.uniform)

    def choice(self, seq):
        return self._t._call("choice", (seq,), self._r.choice)


class _NpRandomFacade:
    def __init__(self, tape: Tape):
        self._t = tape

    def choice(self, a, p=None):
        return self._t._call("np.choice", (a, p), lambda a_, p_: np.random.choice(a_, p=p_))

    def uniform(self, lo, hi):
        return self._t._call("np.uniform", (lo, hi), np.random.uniform)

    def __getattr__(self, k):
        return getattr(np.random, k)


def _centre_from_draw(r, args):
    n = args[1] + 1
    return (int(r) * n) >> 32


def _centre_to_draw(idx, args):
    n = args[1] + 1
    r = -((-int(idx) << 32) // n)          # ceil(idx * 2^32 / n)
    assert (r * n) >> 32 == idx and r < (1 << 32)
    return r


def _index_of(seq, v):
    for i, s in enumerate(seq):
        if s is v:
            return i
    return list(seq).index(v)


def _defect_program(t: Tape, req, y: int, n_coords: int, box):
    """Draws of the y==1 / y==2 branches (datasets.py:267-355 == do_patch/do_scar bodies).
    Generator helper: ``req`` is the first pending request; returns nothing; the caller
    regains control when the reference stops drawing."""
    L, T, R, Bm = box

    def patch_draws(req):
        # np.random.choice([0,1,2], p=[0.7,0.15,0.15])   datasets.py:270/311
        req = yield t.val(req, "variant", want="np.choice")
        variant = int(t.draws["variant"])
        # generate_patch  dataset_generator.py:173-199
        req = yield t.val(req, "area_u", want="uniform")
        req = yield t.val(req, "asp0", want="uniform")
        req = yield t.val(req, "asp1", want="uniform")
        req = yield t.val(req, "asp_pick", want="choice", to_draw=lambda v, a: _index_of(a[0], v),
                          from_draw=lambda s, a: a[0][int(s)])
        req = yield t.val(req, "rect_l", want="randint", check=lambda a: a[0] == 0)
        req = yield t.val(req, "rect_t", want="randint", check=lambda a: a[0] == 0)
        if variant == 2:
            for c in range(3):
                req = yield t.val(req, "rnd_rgb", c, want="randint", check=lambda a: a == (0, 255))
        # similarity branch (data dependent: the reference decides)  datasets.py:295-299
        if req[0] == "np.uniform":
            req = yield t.val(req, "low", want="np.uniform", check=lambda a: a == (0.75, 0.9))
            req = yield t.val(req, "high", want="np.uniform", check=lambda a: a == (1.1, 1.15))
            req = yield t.val(req, "pick1", want="choice", to_draw=lambda v, a: _index_of(a[0], v),
                              from_draw=lambda s, a: a[0][int(s)])
            req = yield t.val(req, "pick2", want="choice", to_draw=lambda v, a: _index_of(a[0], v),
                              from_draw=lambda s, a: a[0][int(s)])
        return req

    def centre(req, i):
        if n_coords >= 2:   # get_random_coordinate  dataset_generator.py:259-265
            req = yield t.val(req, "centre_r", i, want="randint", check=lambda a: a == (0, n_coords - 1),
                              to_draw=_centre_to_draw, from_draw=_centre_from_draw)
        return req

    def fallback(req, i):
        if n_coords == 0:   # check_valid_coordinates_by_container  dataset_generator.py:112-114
            t.draws["fallback_c"][i] = t.draws["fallback_c"][i]
            name, args, real = req
            assert name == "randint" and args == (L, R), f"fallback cx: {name}{args} vs {(L, R)}"
            if t.record:
                v = real(*args)
                t.draws["fallback_c"][i][0] = v
            else:
                v = int(t.draws["fallback_c"][i][0])
            req = yield v
            name, args, real = req
            assert name == "randint" and args == (T, Bm)
            if t.record:
                v = real(*args)
                t.draws["fallback_c"][i][1] = v
            else:
                v = int(t.draws["fallback_c"][i][1])
            req = yield v
        return req

    if y == 1:
        req = yield from centre(req, 0)
        req = yield from patch_draws(req)
        req = yield from fallback(req, 0)
        # rect2poly  dataset_generator.py:65-98
        for side in range(4):
            req = yield t.val(req, "poly_n_side", side, want="randint", check=lambda a: a == (1, 2))
            for j in range(int(t.draws["poly_n_side"][side])):
                req = yield t.val(req, "poly_v", (side, j), want="randint")
    elif y == 2:
        req = yield from patch_draws(req)
        req = yield t.val(req, "scar_k", want="randint", check=lambda a: a == (2, 5))
        req = yield t.val(req, "scar_angle", want="randint", check=lambda a: a == (-45, 45))
        for i in range(int(t.draws["scar_k"])):
            req = yield from centre(req, i)
            req = yield from fallback(req, i)
    return req


def _randint_from_draw(r, args):
    a, b = args
    return a + ((int(r) * (b - a + 1)) >> 32)


def _randint_to_draw(v, args):
    a, b = args
    n = b - a + 1
    r = -((-(int(v) - a) << 32) // n)          # ceil((v - a) * 2^32 / n)
    assert a + ((r * n) >> 32) == v and r < (1 << 32)
    return r


def _line_program(t: Tape, req, n_coords: int, patch_localization: bool):
    """Draws of the y == 3 branch (datasets.py:357-388): side, n index picks, colour, (image-level) split pick."""
    req = yield t.val(req, "line_side", want="choice", to_draw=lambda v, a: _index_of(a[0], v),
                      from_draw=lambda s, a: a[0][int(s)])
    n = 30 if patch_localization else 60
    for i in range(n):
        req = yield t.val(req, "line_r", i, want="randint", check=lambda a, i=i: a[1] == int(n_coords * (i / n)),
                          to_draw=_randint_to_draw, from_draw=_randint_from_draw)
    req = yield t.val(req, "line_rgb", want="choice", to_draw=lambda v, a: _index_of(a[0], v),
                      from_draw=lambda s, a: a[0][int(s)])
    if not patch_localization:
        req = yield t.val(req, "line_k", want="randint", check=lambda a: a == (0, 9))
    return req


def _no_more(req):
    raise AssertionError(f"unexpected extra draw {req[0]}{req[1]}")


@contextlib.contextmanager
def _patched(mods, tape: Tape, torch_params: Dict):
    """Install the tape into the reference modules + torchvision get_params."""
    from torchvision import transforms
    import torch
    rnd = _RandomFacade(tape)
    nprnd = _NpRandomFacade(tape)
    saved = [(m, "random", m.random) for m in (mods.dg, mods.ds, mods.tat)]
    saved += [(m, "np", m.np) for m in (mods.ds, mods.tat)]
    ra, rc, cj = transforms.RandomAffine.get_params, transforms.RandomCrop.get_params, transforms.ColorJitter.get_params
    d, rec = tape.draws, tape.record

    def affine_params(degrees, translate, scale_ranges, shears, img_size):
        if rec:
            out = ra(degrees, translate, scale_ranges, shears, img_size)
            d["warp_angle"], d["warp_scale"] = out[0], out[2]
            return out
        return float(d["warp_angle"]), (0, 0), float(d["warp_scale"]), (0.0, 0.0)

    def crop_params(img, output_size):
        if rec:
            out = rc(img, output_size)
            d["donor_i"], d["donor_j"] = out[0], out[1]
            return out
        return int(d["donor_i"]), int(d["donor_j"]), output_size[0], output_size[1]

    def jitter_params(brightness, contrast, saturation, hue):
        if rec:
            out = cj(brightness, contrast, saturation, hue)
            d["jit_order"] = np.asarray(out[0].tolist())
            d["jit_b"], d["jit_c"], d["jit_s"] = out[1], out[2], out[3]
            return out
        return torch.as_tensor(np.asarray(d["jit_order"], dtype=np.int64)), float(d["jit_b"]), float(d["jit_c"]), float(d["jit_s"]), None

    try:
        for m in (mods.dg, mods.ds, mods.tat):
            m.random = rnd
        for m in (mods.ds, mods.tat):
            m.np = _NpProxy(nprnd)
        transforms.RandomAffine.get_params = staticmethod(affine_params)
        transforms.RandomCrop.get_params = staticmethod(crop_params)
        transforms.ColorJitter.get_params = staticmethod(jitter_params)
        yield
    finally:
        for m, k, v in saved:
            setattr(m, k, v)
        transforms.RandomAffine.get_params = staticmethod(ra)
        transforms.RandomCrop.get_params = staticmethod(rc)
        transforms.ColorJitter.get_params = staticmethod(cj)


# ------------------------------------------------------------------ drivers
class _ListOfOne:
    """images_filenames stand-in: ``Image.open(self.images_filenames[index])``."""

    def __init__(self, path):
        self.path = path
        self.shape = (1,)

    def __getitem__(self, i):
        return self.path


def make_dataset(cfg, original_png: str, donors, fixed_segmentation, transform=None, jitter=True):
    """Build a reference ``PretextTaskDataset`` without running its ``__init__`` I/O
    (datasets.py:166-206 reads ``dataset/<class>/train/good/000.png`` from the CWD);
    the attributes ``__getitem__`` reads are set directly."""
    mods = load_reference()
    ds = mods.ds.PretextTaskDataset.__new__(mods.ds.PretextTaskDataset)
    ds.subject = cfg.subject
    ds.images_filenames = _ListOfOne(original_png)
    ds.imsize = tuple(cfg.imsize)
    ds.transform = transform
    ds.patch_localization = cfg.patch_localization
    ds.patch_size = cfg.patch_size
    ds.patch_area_ratio = tuple(cfg.patch_area_ratio)
    ds.scar_area_ratio = tuple(cfg.scar_area_ratio)
    ds.images_for_cut = donors
    ds.fixed_segmentation = fixed_segmentation
    return ds


def run_getitem(cfg, original_png: str, draws: Dict, donors, fixed_segmentation, n_coords: int,
                effective_label: Optional[int] = None, record: bool = False, with_transform: bool = False):
    """Call the reference's ``PretextTaskDataset.__getitem__`` (datasets.py:209-394) with
    the tape installed.  ``n_coords`` = number of set mask pixels in the window the
    reference will look at (needed to know whether ``get_random_coordinate`` draws);
    ``effective_label`` = label after the coverage rule (defaults to draws['label']).
    Returns ``(x, y, original_tensor)`` exactly as the reference returns them
    (``x`` is a PIL image when ``with_transform`` is False)."""
    mods = load_reference()
    from torchvision import transforms
    transform = None
    if with_transform:
        transform = transforms.Compose([transforms.ToTensor(),
                                        transforms.Normalize((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))])
    ds = make_dataset(cfg, original_png, donors, fixed_segmentation, transform)
    box1 = cfg.box_poly
    box2 = cfg.box_scar

    def program(t: Tape):
        req = yield
        # datasets.py:215 -- HEAD draws randint(0,3); the 3-class contract never answers 3
        req = yield t.val(req, "label", want="randint", check=lambda a: a == (0, 3))
        if cfg.is_texture:   # datasets.py:226
            req = yield t.val(req, "donor_idx", want="choice", to_draw=lambda v, a: _index_of(a[0], v),
                              from_draw=lambda s, a: a[0][int(s)])
        if cfg.patch_localization:   # datasets.py:249-250
            req = yield t.val(req, "crop_left", want="randint", check=lambda a: a[0] == 0)
            req = yield t.val(req, "crop_top", want="randint", check=lambda a: a[0] == 0)
        y = int(t.draws["label"]) if effective_label is None else effective_label
        if y in (1, 2):
            req = yield from _defect_program(t, req, y, n_coords, box1 if y == 1 else box2)
        elif y == 3:
            req = yield from _line_program(t, req, n_coords, cfg.patch_localization)
        _no_more(req)

    tape = Tape(program, draws, record)
    with _patched(mods, tape, {}):
        if not cfg.jitter:
            saved_j = mods.ds.CPP.jitter_transforms
            mods.ds.CPP.jitter_transforms = lambda im: im
        try:
            out = ds[0]
        finally:
            if not cfg.jitter:
                mods.ds.CPP.jitter_transforms = saved_j
    return out


def run_script_fn(which: str, img, image_for_cutting, segmentation, patch_loc: bool, draws: Dict,
                  n_coords: int, box, record: bool = False):
    """Call the test script's ``do_patch`` / ``do_scar``
    (test_artificial_transformations.py:20-122) with the tape installed."""
    mods = load_reference()
    fn = {"do_patch": mods.tat.do_patch, "do_scar": mods.tat.do_scar}[which]

    def program(t: Tape):
        req = yield
        if which == "do_patch":
            # test_artificial_transformations.py:23  k = random.randint(1,2)  (unused)
            name, args, real = req
            assert name == "randint" and args == (1, 2)
            req = yield (real(*args) if t.record else 1)
        req = yield from _defect_program(t, req, 1 if which == "do_patch" else 2, n_coords, box)
        _no_more(req)

    tape = Tape(program, draws, record)
    with _patched(mods, tape, {}):
        return fn(img, image_for_cutting=image_for_cutting, segmentation=segmentation, patch_loc=patch_loc)


def reference_obj_mask(image):
    """The reference's ``obj_mask`` body (dataset_generator.py:27-39) executed with the
    restated ``canny`` / ``label`` standing in for the absent skimage."""
    return load_reference().dg.obj_mask(image)
