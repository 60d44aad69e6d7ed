"""The reference-facing host API (``ssad_b200.dataset_generator`` / ``ssad_b200.datasets``): same names and argument
meaning as ``src/self_supervised/dataset_generator.py``, ``datasets.py`` and the exerciser script's do_patch / do_scar.

CPU part: host integer logic against the oracle, file-list logic of the datamodule.  GPU part (``-m gpu``): the
function-level entry points against the golden vectors made from the reference's own do_patch / do_scar, against
Pillow, and the batched dataset against the reference's ``__getitem__`` goldens."""
import os

import numpy as np
import pytest

import scenarios as S
from golden import make_golden as G

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


# ---------------------------------------------------------------------------------------------- CPU
def test_container_and_clamp_match_oracle():
    from oracle import generator as og
    from ssad_b200 import dataset_generator as dg
    rng = np.random.default_rng(0)
    for _ in range(3000):
        W = int(rng.choice([32, 64, 256, 320, 1024]))
        f = float(rng.choice([1, 1.75, 2]))
        size = (int(rng.integers(2, W + 40)), int(rng.integers(2, W + 40)))
        c = (int(rng.integers(0, W)), int(rng.integers(0, W)))
        box = dg.Container((W, W), f)
        ref = og.Container((W, W), f)
        assert (box.left, box.top, box.right, box.bottom) == (ref.left, ref.top, ref.right, ref.bottom)
        assert dg.check_valid_coordinates_by_container((W, W), size, c, f) == \
            tuple(og.check_valid_coordinates_by_container((W, W), size, c, f, None))


def test_clamp_draws_centre_inside_container_when_none():
    from ssad_b200 import dataset_generator as dg
    rng = np.random.default_rng(1)
    box = dg.Container((256, 256), 2)
    for _ in range(200):
        l, t = dg.check_valid_coordinates_by_container((256, 256), (10, 6), None, 2, rng=rng)
        assert box.left <= l <= box.right and box.top <= t <= box.bottom


def test_get_random_coordinate_contract():
    from ssad_b200 import dataset_generator as dg
    assert dg.get_random_coordinate(np.zeros((0, 2))) is None          # dataset_generator.py:260-261
    one = np.array([[3, 4]])
    assert tuple(dg.get_random_coordinate(one)) == (3, 4)
    many = np.arange(20).reshape(10, 2)
    assert any(tuple(dg.get_random_coordinate(many, np.random.default_rng(s))) != (0, 1) for s in range(8))


def test_duplicate_filenames_semantics():
    from ssad_b200 import datasets as ds
    names = np.array(["a", "b", "c"])
    out = ds.duplicate_filenames(names, 10)
    assert len(out) == 12 and list(out[:6]) == ["a", "b", "c", "a", "b", "c"]      # whole copies (functional.py:64-68)
    assert len(ds.duplicate_filenames(names, 2)) == 3


def _png_tree(tmp_path, subject="bottle", n_train=10, n_test=4, size=48):
    from PIL import Image
    from ssad_b200 import synth
    for cls in (subject, "carpet"):
        for split, n in (("train", n_train), ("test", n_test)):
            d = tmp_path / "dataset" / cls / split / "good"
            d.mkdir(parents=True)
            imgs = synth.object_images(n, size, size, 11 + len(cls))
            for i in range(n):
                Image.fromarray(imgs[i]).save(d / f"{i:03d}.png")
    return str(tmp_path / "dataset" / subject)


def test_datamodule_file_lists(tmp_path):
    from ssad_b200 import datasets as ds
    root = _png_tree(tmp_path)
    dm = ds.PretextTaskDatamodule("bottle", root, imsize=(32, 32), batch_size=4, min_dataset_length=30, seed=0)
    assert len(dm.train_images_filenames) >= 30 and len(dm.val_images_filenames) >= 30 and len(dm.test_images_filenames) >= 30
    assert len(set(dm.train_images_filenames)) == 8 and len(set(dm.val_images_filenames)) == 2      # sklearn 0.2 split of 10
    assert not set(dm.train_images_filenames) & set(dm.val_images_filenames)
    dm2 = ds.PretextTaskDatamodule("bottle", root, imsize=(32, 32), duplication=False)
    assert len(dm2.train_images_filenames) == 8


# ---------------------------------------------------------------------------------------------- GPU
@pytest.mark.gpu
@pytest.mark.parametrize("which", ["do_patch", "do_scar"])
@pytest.mark.parametrize("mode", ["image", "patch"])
def test_do_patch_do_scar_match_reference_golden(which, mode):
    """ssad_b200.dataset_generator.do_patch / do_scar on the parameters the golden run used == the reference's own
    functions (test_artificial_transformations.py:20-122), bit for bit."""
    from PIL import Image
    import ssad_b200
    from ssad_b200 import dataset_generator as dg
    g = np.load(os.path.join(GOLD, "script_fns.npz"))
    sc = S.build("image_level", 8, seed=7)
    assert bytes(g["sha"]).hex() == G.scenario_sha(sc)
    seg = S.mask_to_pil(sc.mask)
    key = f"{which}_{mode}"
    draws = {k[len(key) + 3:]: g[k] for k in g.files if k.startswith(key + "_d_")}
    cfg = dg._script_config(sc.cfg.imsize, mode == "patch")
    pb = ssad_b200.params_from_draws(cfg, draws)
    rec_gold = np.frombuffer(g[key + "_rec"].tobytes(), dtype=pb.records.dtype)
    for f in ("rect", "poly_xy", "poly_n", "scar_fix", "scar_size", "scar_k", "src_variant", "centre_r", "bright_f"):
        assert np.array_equal(pb.records[f], rec_gold[f]), f        # (the golden records also carry an unused pre-warp)
    fn = dg.do_patch if which == "do_patch" else dg.do_scar
    for i in range(len(sc.images)):
        img = Image.fromarray(sc.images[i])
        x = fn(img, img, seg, mode == "patch", params=pb.slice(i, i + 1))
        assert isinstance(x, Image.Image)
        assert np.array_equal(np.array(x), g[key + "_x"][i]), (key, i)


@pytest.mark.gpu
def test_do_patch_samples_its_own_parameters():
    from PIL import Image
    from ssad_b200 import dataset_generator as dg
    sc = S.build("image_level", 4, seed=3)
    img = Image.fromarray(sc.images[0])
    a = np.array(dg.do_patch(img, img, S.mask_to_pil(sc.mask), False, rng=np.random.default_rng(5)))
    b = np.array(dg.do_patch(img, img, S.mask_to_pil(sc.mask), False, rng=np.random.default_rng(5)))
    assert np.array_equal(a, b)                                   # same generator state -> same defect
    changed = (a != sc.images[0]).any(axis=2)
    assert 0 < changed.sum() < 0.08 * changed.size                 # area_ratio (0.03, 0.07): a patch, not the image


@pytest.mark.gpu
def test_rect2poly_matches_pillow():
    from PIL import Image, ImageDraw
    from ssad_b200 import dataset_generator as dg
    rng = np.random.default_rng(2)
    for _ in range(40):
        pw, ph = int(rng.integers(2, 130)), int(rng.integers(2, 130))
        ns, v = dg._redraw_vertices(pw, ph, rng)
        from ssad_b200.params import polygon_vertices
        pts = polygon_vertices(pw, ph, ns, v)
        ref = Image.new("RGBA", (pw, ph), (0, 0, 0, 0))
        ImageDraw.Draw(ref).polygon([tuple(int(c) for c in p) for p in pts], fill="white")     # dataset_generator.py:99-100
        got = dg.rect2poly(Image.new("RGB", (pw, ph)), points=pts)
        assert got.mode == "RGBA" and np.array_equal(np.array(got), np.array(ref)), (pw, ph, pts)


@pytest.mark.gpu
def test_paste_patch_matches_pillow():
    from PIL import Image
    from ssad_b200 import dataset_generator as dg
    rng = np.random.default_rng(4)
    img = Image.fromarray(rng.integers(0, 256, (40, 56, 3), dtype=np.uint8))
    for _ in range(30):
        pw, ph = int(rng.integers(2, 30)), int(rng.integers(2, 30))
        patch = Image.fromarray(rng.integers(0, 256, (ph, pw, 3), dtype=np.uint8))
        alpha = (rng.random((ph, pw)) < 0.6).astype(np.uint8) * 255
        rgba = Image.fromarray(np.dstack([np.array(patch), alpha]), mode="RGBA")
        at = (int(rng.integers(-20, 60)), int(rng.integers(-20, 45)))     # partly or wholly outside: Pillow clips
        for p, m in ((patch, None), (patch, rgba), (rgba, rgba)):
            ref = img.copy()
            ref.paste(p, at, mask=m)                                 # dataset_generator.py:273-274
            assert np.array_equal(np.array(dg.paste_patch(img, p, at, m)), np.array(ref)), (at, pw, ph)


@pytest.mark.gpu
def test_paste_patch_soft_mask_blends_like_pillow():
    """A mask with values between 0 and 255 (the reference's commented-out GaussianBlur mask is one): Paste.c blends."""
    from PIL import Image
    from ssad_b200 import dataset_generator as dg
    rng = np.random.default_rng(8)
    img = Image.fromarray(rng.integers(0, 256, (48, 48, 3), dtype=np.uint8))
    patch = Image.fromarray(rng.integers(0, 256, (20, 31, 3), dtype=np.uint8))
    soft = Image.fromarray(rng.integers(0, 256, (20, 31), dtype=np.uint8), mode="L")
    ref = img.copy()
    ref.paste(patch, (9, -4), mask=soft)
    assert np.array_equal(np.array(dg.paste_patch(img, patch, (9, -4), soft)), np.array(ref))


@pytest.mark.gpu
def test_check_color_similarity_matches_reference_formula():
    """dataset_generator.py:147-159: numpy mean / 255 per channel, sklearn cosine_similarity."""
    from sklearn.metrics.pairwise import cosine_similarity
    from ssad_b200 import dataset_generator as dg
    rng = np.random.default_rng(12)
    for _ in range(20):
        a = rng.integers(0, 256, (int(rng.integers(2, 90)), int(rng.integers(2, 90)), 3), dtype=np.uint8)
        b = rng.integers(0, int(rng.integers(2, 256)), (int(rng.integers(2, 40)), int(rng.integers(2, 40)), 3), dtype=np.uint8)
        v1 = (a.mean(axis=(0, 1)) / 255)[None, :]
        v2 = (b.mean(axis=(0, 1)) / 255)[None, :]
        want = float(cosine_similarity(v1, v2).squeeze())
        assert abs(dg.check_color_similarity(a, b) - want) < 1e-12


@pytest.mark.gpu
def test_obj_mask_pil_in_pil_out():
    from PIL import Image
    from oracle import objmask as om
    from ssad_b200 import dataset_generator as dg, synth
    img = Image.fromarray(synth.screw_images(1, 96, 96, 3)[0])
    got = dg.obj_mask(img)
    ref = om.obj_mask(img)
    assert got.mode == "RGB" and np.array_equal(np.array(got), np.array(ref))


@pytest.mark.gpu
@pytest.mark.parametrize("name", ["image_level", "patch_level", "texture", "non_fixed"])
def test_dataset_batch_matches_reference_golden(name):
    """PretextTaskDataset.get_batch with the golden run's parameters == the reference's __getitem__ outputs."""
    import torch
    from ssad_b200 import datasets as ds
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    sc = S.build(name, G.BATCH[name])
    assert bytes(g["sha"]).hex() == G.scenario_sha(sc)
    fixed = sc.mask if (sc.mask is not None and sc.mask.ndim == 2) else None
    common = dict(imsize=sc.cfg.imsize, patch_localization=sc.cfg.patch_localization, patch_size=sc.cfg.patch_size,
                  images=sc.images, donors=sc.donors, fixed_segmentation=fixed, dataset_dir="/nonexistent/")
    dset = ds.PretextTaskDataset(sc.cfg.subject, transform=True, **common)
    x, y, original = dset.get_batch(np.arange(len(sc.images)), params=sc.params)
    assert np.array_equal(y.cpu().numpy(), g["y"])
    assert np.array_equal(x.cpu().numpy()[:2], g["xn"])                          # Normalize(ToTensor(x))
    assert torch.equal(original.cpu(), torch.from_numpy(sc.images).permute(0, 3, 1, 2).float().div(255))   # ToTensor(original)
    raw = ds.PretextTaskDataset(sc.cfg.subject, transform=False, **common)
    xu, yu, _ = raw.get_batch(np.arange(len(sc.images)), params=sc.params)
    assert np.array_equal(xu.cpu().numpy(), g["x"])
    x0, y0, o0 = dset[0]
    assert tuple(x0.shape) == (3,) + tuple(x.shape[2:]) and int(y0) in (0, 1, 2) and tuple(o0.shape) == tuple(original.shape[1:])


@pytest.mark.gpu
def test_datamodule_yields_training_batches(tmp_path):
    import torch
    from ssad_b200 import datasets as ds
    root = _png_tree(tmp_path, size=64)
    dm = ds.PretextTaskDatamodule("bottle", root, imsize=(64, 64), batch_size=8, min_dataset_length=40, seed=0)
    dm.setup()
    n = 0
    for x, y, original in dm.train_dataloader():
        assert x.shape == (8, 3, 64, 64) and x.dtype == torch.float32 and x.is_cuda
        assert y.shape == (8,) and y.dtype == torch.int64 and int(y.min()) >= 0 and int(y.max()) <= 2
        assert original.shape == (8, 3, 64, 64) and float(original.max()) <= 1.0
        n += 1
    assert n == len(dm.train_dataloader()) == 5                      # 40 items, drop_last
    assert len(dm.test_dataloader()) == 5
    xt, yt, _ = next(iter(dm.test_dataloader()))
    assert xt.shape[0] == 8


@pytest.mark.gpu
@pytest.mark.parametrize("patch_loc", [False, True])
def test_do_swirl_matches_oracle(patch_loc):
    """do_swirl (test_artificial_transformations.py:125-164) = label 3 on one image, against the oracle's restatement
    (pinned by the label-3 goldens made from the reference's own __getitem__)."""
    from PIL import Image
    from oracle import generator as og
    from ssad_b200 import dataset_generator as dg, sample_params
    sc = S.build("image_level", 6, seed=3)
    seg = S.mask_to_pil(sc.mask)
    cfg = dg._script_config(sc.cfg.imsize, patch_loc)
    for i in range(6):
        pb = sample_params(cfg, 1, np.random.default_rng(70 + i), labels=np.full(1, 3))
        img = Image.fromarray(sc.images[i])
        got = dg.do_swirl(img, img, seg, patch_loc, params=pb)
        want = og.do_swirl(img, img, seg, patch_loc, pb.image_draws(0))
        assert np.array_equal(np.array(got), np.array(want)), (patch_loc, i)
        assert not np.array_equal(np.array(got), sc.images[i])


# ---------------------------------------------------------------------------------------------- MVTec evaluation loader
def _write_mvtec_tree(root, n_train=6, n_test=3, size=(40, 32)):
    """A tiny MVTec-shaped class directory with PNG files: train/good, test/{good,crack}, ground_truth/crack."""
    from PIL import Image
    rng = np.random.default_rng(2)
    for sub, n in (("train/good", n_train), ("test/good", n_test), ("test/crack", n_test)):
        os.makedirs(os.path.join(root, sub), exist_ok=True)
        for i in range(n):
            Image.fromarray(rng.integers(0, 256, (size[1], size[0], 3), dtype=np.uint8)).save(os.path.join(root, sub, f"{i:03d}.png"))
    os.makedirs(os.path.join(root, "ground_truth/crack"), exist_ok=True)
    for i in range(n_test):
        m = (rng.random((size[1], size[0])) < 0.3).astype(np.uint8) * 255
        Image.fromarray(m, mode="L").save(os.path.join(root, "ground_truth/crack", f"{i:03d}_mask.png"))


def test_mvtec_filename_helpers_match_reference_rules(tmp_path):
    from ssad_b200 import datasets as D
    root = str(tmp_path / "bottle") + "/"
    _write_mvtec_tree(root)
    names = D.get_test_data_filenames(root + "test/")
    assert [n.rsplit("/", 2)[1] for n in names] == ["crack"] * 3 + ["good"] * 3          # class by class, sorted
    assert D.get_ground_truth_filename(names[0], root + "ground_truth/") == root + "ground_truth/crack/000_mask.png"
    assert D.get_ground_truth_filename(names[-1], root + "ground_truth/") is None          # functional.py:48-49
    assert not D.get_ground_truth(None, (8, 6)).any() and D.get_ground_truth(None, (8, 6)).shape == (6, 8)


@pytest.mark.gpu
def test_mvtec_datamodule_matches_torchvision_items(tmp_path):
    """MVTecDataset.__getitem__ (datasets.py:67-81): x = Normalize(ToTensor(resized)), ToTensor(gt), ToTensor(original)."""
    import torch
    from PIL import Image
    from torchvision import transforms
    from ssad_b200 import datasets as D
    root = str(tmp_path / "bottle") + "/"
    _write_mvtec_tree(root)
    dm = D.MVTecDatamodule(root, imsize=(24, 16), batch_size=4, seed=0)
    dm.setup()
    tf = transforms.Compose([transforms.ToTensor(), transforms.Normalize((0.485, 0.456, 0.406), (0.229, 0.224, 0.225))])
    seen = 0
    for x, gt, original in dm.test_dataloader():
        for k in range(x.shape[0]):
            f = dm.test_images_filenames[seen]
            img = Image.open(f).resize((24, 16)).convert("RGB")
            g = D.get_ground_truth_filename(f, root + "ground_truth/")
            gt_ref = Image.open(g).resize((24, 16)).convert("1") if g else Image.new(mode="1", size=(24, 16))
            assert torch.equal(x[k].cpu(), tf(img))
            assert torch.equal(original[k].cpu(), transforms.ToTensor()(img))
            assert torch.equal(gt[k].cpu(), transforms.ToTensor()(gt_ref))
            seen += 1
    assert seen == len(dm.test_images_filenames) == 6
    assert len(dm.train_dataset) + len(dm.val_dataset) == 6
