"""Batched, device-resident mirror of the reference's data layer (``src/self_supervised/datasets.py``):
``PretextTaskDataset`` (:166-398) and ``PretextTaskDatamodule`` (:401-533), same constructor arguments.

The reference produces one item per ``__getitem__`` call inside 8 DataLoader worker processes; here the decoded
images live on the GPU as uint8 and a whole batch goes through one ``ssad_generate`` call.  A loader yields the tuples
``training_step`` consumes (models.py:256-261): ``(x f32|bf16 [B,3,h,w] normalised, y int64 [B], original f32 [B,3,H,W])``
-- already on the device.  Randomness: one ``numpy.random.Generator`` per dataset (seeded), parameters drawn on the
host by ``params.sample_params`` with the reference's distributions; pass ``params=`` to replay explicit draws.
"""
from __future__ import annotations

import ctypes as C
import glob
import os
from typing import Iterator, Optional, Sequence, Tuple

import numpy as np
import torch

from . import _native as N
from . import constants
from .config import CPP, GeneratorConfig
from .generator import BatchGenerator, obj_mask_batch, unpack_mask_bits
from .params import ParamBatch, sample_params


def get_filenames(main_path: str) -> np.ndarray:
    """functional.py:32-35: sorted ``*.png`` directly under ``main_path``."""
    return np.array(sorted(f.replace("\\", "/") for f in glob.glob(main_path + "*.png")))


def duplicate_filenames(filenames: np.ndarray, baseline: int = 2000) -> np.ndarray:
    """functional.py:64-68: whole copies of the list are appended until it holds at least ``baseline`` names."""
    filenames = np.asarray(filenames)
    out = filenames.copy()
    while out.shape[0] < baseline and filenames.shape[0] > 0:
        out = np.concatenate([out, filenames])
    return out


def load_image(filename: str, imsize: Tuple[int, int]) -> np.ndarray:
    """datasets.py:211-213: ``Image.open(f).resize(imsize).convert('RGB')`` (PIL's default bicubic filter).  Host
    side: decoding is upstream of the hot path; each distinct file is decoded once."""
    from PIL import Image
    return np.asarray(Image.open(filename).resize(tuple(imsize)).convert("RGB"))


def to_tensor(images_u8: torch.Tensor) -> torch.Tensor:
    """``ToTensor(original)`` (datasets.py:394) for a batch: uint8 [B,H,W,3] on the device -> f32 [B,3,H,W] / 255."""
    lib = N.load()
    B, H, W = int(images_u8.shape[0]), int(images_u8.shape[1]), int(images_u8.shape[2])
    out = torch.empty((B, 3, H, W), dtype=torch.float32, device=images_u8.device)
    stream = torch.cuda.current_stream(images_u8.device).cuda_stream
    N.check(lib.ssad_to_tensor(B, H, W, C.c_void_p(images_u8.data_ptr()), int(images_u8.stride(1)), int(images_u8.stride(0)),
                               C.c_void_p(out.data_ptr()), C.c_void_p(stream)), "ssad_to_tensor")
    return out


class PretextTaskDataset:
    """datasets.py:166-398.  ``images_filenames`` as in the reference (duplicates allowed: each distinct file is
    decoded once), or ``images=`` uint8 [N,H,W,3] already decoded.  Indexing with an int returns one item like the
    reference's ``__getitem__``; ``get_batch(indices)`` is the fast path."""

    def __init__(self, subject: str, images_filenames: Optional[Sequence[str]] = None, imsize: tuple = (256, 256),
                 transform=True, patch_localization: bool = False, patch_size: int = 64, *,
                 images: Optional[np.ndarray] = None, donors: Optional[np.ndarray] = None,
                 fixed_segmentation: Optional[np.ndarray] = None, dataset_dir: str = "dataset/", device="cuda",
                 out_dtype: torch.dtype = torch.float32, seed: Optional[int] = None, num_labels: int = 3,
                 return_original: bool = True) -> None:
        self.subject = subject
        self.imsize = tuple(imsize)
        self.transform = transform
        self.patch_localization = patch_localization
        self.patch_size = patch_size
        self.device = torch.device(device)
        self.out_dtype = out_dtype
        self.return_original = return_original
        self.rng = np.random.default_rng(seed)
        W, H = self.imsize
        # ---- the decoded originals, resident on the device; index[i] = row of item i
        if images is not None:
            images = np.ascontiguousarray(images, dtype=np.uint8)
            if images.ndim != 4 or images.shape[1:] != (H, W, 3):
                raise ValueError(f"images must be uint8 [N, {H}, {W}, 3]")
            self.images_filenames = np.arange(images.shape[0])
            self.index = np.arange(images.shape[0])
            uniq = images
        else:
            self.images_filenames = np.asarray(images_filenames)
            names, self.index = np.unique(self.images_filenames, return_inverse=True)
            uniq = np.stack([load_image(f, self.imsize) for f in names]) if len(names) else np.zeros((0, H, W, 3), np.uint8)
        self.images = torch.from_numpy(uniq).to(self.device)
        # ---- donor bank: image 000 of every class under dataset_dir (datasets.py:189-193); textures need it
        is_texture = subject in constants.TEXTURES()
        if donors is None and os.path.isdir(dataset_dir):
            classes = sorted(n for n in os.listdir(dataset_dir) if os.path.isdir(os.path.join(dataset_dir, n)))
            files = [os.path.join(dataset_dir, c, "train/good/000.png") for c in classes]
            if files and all(os.path.exists(f) for f in files) and is_texture:
                donors = np.stack([load_image(f, self.imsize) for f in files])
        n_donors = 0 if donors is None else int(np.asarray(donors).shape[0])
        self.cfg = GeneratorConfig(subject=subject, imsize=self.imsize, patch_localization=patch_localization,
                                   patch_size=patch_size, n_donors=n_donors, num_labels=num_labels)
        # ---- the mask policy (datasets.py:195-206, 232-235): white for textures, obj_mask(image 000) for fixed-pose
        #      classes (cached), per image for hazelnut / screw / metal_nut
        mask = None
        if not is_texture and not self.cfg.is_non_fixed:
            if fixed_segmentation is not None:
                mask = np.asarray(fixed_segmentation).astype(bool)
                if mask.ndim == 3:
                    mask = mask[..., 0]
            else:
                f000 = os.path.join(dataset_dir, subject, "train/good/000.png")
                first = torch.from_numpy(load_image(f000, self.imsize)[None]).to(self.device) if os.path.exists(f000) else self.images[:1]
                # cable: the reference averages SLIC super-pixels first (datasets.py:201-205, skimage.segmentation.slic
                # -- a one-time host step outside the hot path, not built): without fixed_segmentation= the mask is the
                # obj_mask of the raw image, which differs from the reference's -- say so instead of doing it silently
                if subject == "cable":
                    import warnings
                    warnings.warn("cable: the reference builds its fixed mask from SLIC super-pixels (datasets.py:201-205); "
                                  "pass fixed_segmentation= to reproduce it -- using obj_mask of the raw image 000", stacklevel=2)
                mask = unpack_mask_bits(obj_mask_batch(first).cpu().numpy(), W)[0]
        self.fixed_segmentation = mask
        self._src_masks = None
        self.gen = BatchGenerator(self.cfg, self.device, mask=mask, donors=donors if is_texture else None)

    def __len__(self) -> int:
        return int(self.index.shape[0])

    def sample(self, batch: int, labels: Optional[np.ndarray] = None) -> ParamBatch:
        return sample_params(self.cfg, batch, self.rng, labels=labels)

    def source_masks(self) -> Optional[torch.Tensor]:
        """Non-fixed classes: ``obj_mask(original)`` (datasets.py:232-233) of every resident image, built on first use.
        The reference recomputes it on every ``__getitem__``; it depends on the un-warped original only, so one mask per
        distinct file is the same result."""
        if not self.cfg.is_non_fixed:
            return None
        if self._src_masks is None:
            self._src_masks = obj_mask_batch(self.images)
        return self._src_masks

    def get_batch(self, indices: Sequence[int], params: Optional[ParamBatch] = None):
        """items ``indices`` -> (x [B,3,h,w], y [B] int64, original [B,3,H,W] f32 | None), all on the device.  The
        resident images are never gathered: every record carries the row of its source image (``src_index``)."""
        rows = self.index[np.asarray(indices)]
        if params is None:
            params = self.sample(len(rows))
        masks = self.source_masks()
        if self.transform:
            out = self.gen.generate(self.images, params, masks=masks, want_u8=False, norm_dtype=self.out_dtype, src_index=rows)
            x = out["x"]
        else:       # no transform in the reference = the PIL image; here the uint8 tensor [B,h,w,3]
            out = self.gen.generate(self.images, params, masks=masks, want_u8=True, src_index=rows)
            x = out["x_u8"]
        original = None
        if self.return_original:                    # ToTensor(original), datasets.py:394 (unused by training_step, models.py:257)
            original = to_tensor(self.images.index_select(0, torch.as_tensor(rows, device=self.device)))
        return x, out["y"], original

    def __getitem__(self, index: int):
        x, y, original = self.get_batch([int(index)])
        return x[0], y[0], (original[0] if original is not None else None)


class BatchLoader:
    """What ``DataLoader(dataset, batch_size, shuffle, drop_last)`` yields in the reference, produced batch-wise."""

    def __init__(self, dataset: PretextTaskDataset, batch_size: int, shuffle: bool, drop_last: bool):
        self.dataset, self.batch_size, self.shuffle, self.drop_last = dataset, int(batch_size), shuffle, drop_last

    def __len__(self) -> int:
        n = len(self.dataset)
        return n // self.batch_size if self.drop_last else (n + self.batch_size - 1) // self.batch_size

    POOL = 2048        # records drawn per call of the sampler (its ~60 NumPy calls cost the same for 96 records as for 2048)

    def __iter__(self) -> Iterator:
        """The parameter records are drawn on a background thread, a pool of several batches at a time, while the
        batches of the previous pool are generated and consumed (the sampler is the host-side
        limiter; drawn batch by batch its fixed cost would dominate at the reference's batch size of 96)."""
        import threading
        n = len(self.dataset)
        ds = self.dataset
        order = ds.rng.permutation(n) if self.shuffle else np.arange(n)
        sels = [order[b * self.batch_size:(b + 1) * self.batch_size] for b in range(len(self))]
        per_pool = max(1, self.POOL // max(self.batch_size, 1))
        pools = [sels[i:i + per_pool] for i in range(0, len(sels), per_pool)]
        box = {}

        def draw(k):
            box[k] = ds.sample(sum(len(s_) for s_ in pools[k]))
        th = None
        if pools:
            th = threading.Thread(target=draw, args=(0,), daemon=True)
            th.start()
        for k, pool in enumerate(pools):
            th.join()
            params = box.pop(k)
            if k + 1 < len(pools):
                th = threading.Thread(target=draw, args=(k + 1,), daemon=True)
                th.start()
            lo = 0
            for sel in pool:
                yield ds.get_batch(sel, params.slice(lo, lo + len(sel)))
                lo += len(sel)


class PretextTaskDatamodule:
    """datasets.py:401-533 with the same arguments; the loaders are ``BatchLoader``s over device-resident datasets.
    Not a LightningDataModule (pytorch_lightning is a dependency of the reference's trainer, not of this path); it has
    the same ``setup`` / ``*_dataloader`` methods, so ``trainer.fit(model, train_dataloaders=dm.train_dataloader())``
    works unchanged."""

    def __init__(self, subject: str, root_dir: str, imsize: tuple = (256, 256), batch_size: int = 32,
                 train_val_split: float = 0.2, seed: int = 0, min_dataset_length: int = 1000, duplication: bool = True,
                 patch_localization: bool = False, patch_size: int = 64, *, device="cuda",
                 out_dtype: torch.dtype = torch.float32, dataset_dir: Optional[str] = None, return_original: bool = True):
        self.subject, self.imsize, self.batch_size = subject, tuple(imsize), batch_size
        self.train_val_split, self.seed = train_val_split, seed
        self.min_dataset_length, self.duplication = min_dataset_length, duplication
        self.patch_localization, self.patch_size = patch_localization, patch_size
        self.device, self.out_dtype, self.return_original = device, out_dtype, return_original
        self.root_dir_train = root_dir + "/train/good/"
        self.root_dir_test = root_dir + "/test/good/"
        self.dataset_dir = dataset_dir if dataset_dir is not None else os.path.dirname(os.path.normpath(root_dir)) + "/"
        self.prepare_filenames()

    def prepare_filenames(self) -> None:
        """datasets.py:441-466: sklearn split of train/good, test/good, duplication to ``min_dataset_length``."""
        from sklearn.model_selection import train_test_split
        names = get_filenames(self.root_dir_train)
        train, val = train_test_split(names, test_size=self.train_val_split, random_state=self.seed)
        test = get_filenames(self.root_dir_test)
        if self.duplication:
            train = duplicate_filenames(train, self.min_dataset_length)
            val = duplicate_filenames(val, self.min_dataset_length)
            test = duplicate_filenames(test, self.min_dataset_length)
        self.train_images_filenames, self.val_images_filenames, self.test_images_filenames = train, val, test

    def prepare_data(self) -> None:
        pass

    def _dataset(self, filenames, salt: int) -> PretextTaskDataset:
        return PretextTaskDataset(self.subject, filenames, imsize=self.imsize, transform=True,
                                  patch_localization=self.patch_localization, patch_size=self.patch_size,
                                  dataset_dir=self.dataset_dir, device=self.device, out_dtype=self.out_dtype,
                                  seed=None if self.seed is None else self.seed + salt, return_original=self.return_original)

    def setup(self, stage: Optional[str] = None) -> None:
        if stage == "fit" or stage is None:
            # as in the reference (datasets.py:475-489) the training set is built from the VALIDATION file list and
            # vice versa; kept so that the same files reach the same loaders
            self.train_dataset = self._dataset(self.val_images_filenames, 1)
            self.val_dataset = self._dataset(self.train_images_filenames, 2)
        if stage in ("test", "predict") or stage is None:
            self.test_dataset = self._dataset(self.test_images_filenames, 3)

    def train_dataloader(self) -> BatchLoader:
        return BatchLoader(self.train_dataset, self.batch_size, shuffle=True, drop_last=True)

    def val_dataloader(self) -> BatchLoader:
        return BatchLoader(self.val_dataset, self.batch_size, shuffle=False, drop_last=True)

    def test_dataloader(self) -> BatchLoader:
        return BatchLoader(self.test_dataset, self.batch_size, shuffle=False, drop_last=False)

    def predict_dataloader(self) -> BatchLoader:
        return BatchLoader(self.test_dataset, self.batch_size, shuffle=False, drop_last=False)


# ---------------------------------------------------------------------- MVTec evaluation data (datasets.py:50-163)
def get_subdirectories(main_path: str) -> np.ndarray:
    """functional.py:38-41."""
    return np.array(sorted(n for n in os.listdir(main_path) if os.path.isdir(main_path + "/" + n)), dtype=str)


def get_test_data_filenames(main_path: str) -> np.ndarray:
    """functional.py:54-61: the ``*.png`` of every defect class under ``<root>/test/``, class by class."""
    out = np.empty(0)
    for defection in get_subdirectories(main_path):
        out = np.concatenate([out, get_filenames(main_path + defection + "/")])
    return out


def get_ground_truth_filename(test_filename: str, ground_truth_dir: str) -> Optional[str]:
    """functional.py:44-51: ``.../test/<defect>/<name>.png`` -> ``<gt dir><defect>/<name>_mask.png``; None for 'good'."""
    parts = test_filename.rsplit("/", 2)
    defection, image_name = parts[1], parts[2]
    if defection == "good":
        return None
    stem = image_name.split(".")
    return ground_truth_dir + defection + "/" + stem[0] + "_mask" + "." + stem[1]


def get_ground_truth(filename: Optional[str] = None, imsize=(256, 256)) -> np.ndarray:
    """functional.py:20-24: ``Image.open(f).resize(imsize).convert('1')`` (all-zero when there is no mask) -> bool [H, W]."""
    from PIL import Image
    if filename:
        return np.asarray(Image.open(filename).resize(tuple(imsize)).convert("1"))
    return np.zeros((imsize[1], imsize[0]), dtype=bool)


class MVTecDataset:
    """datasets.py:50-84: real test images with their ground-truth masks, no augmentation.  Every file is decoded once
    on the host; images (uint8) and masks live on the device.  ``get_batch(indices)`` / iteration yield the reference's
    item tuple batch-wise: ``(x, gt f32 [B,1,H,W], original f32 [B,3,H,W])`` with ``x`` = ``Normalize(ToTensor(...))``
    when ``transform`` is set (the datamodule's transform, datasets.py:100-103), else the uint8 images."""

    def __init__(self, dataset_dir: str, images_filenames, imsize: tuple = (256, 256), transform=True,
                 patch_level: bool = False, *, device="cuda") -> None:
        self.dataset_dir, self.imsize, self.transform, self.patch_level = dataset_dir, tuple(imsize), transform, patch_level
        self.images_filenames = np.asarray(images_filenames)
        self.device = torch.device(device)
        W, H = self.imsize
        n = len(self.images_filenames)
        imgs = np.zeros((n, H, W, 3), dtype=np.uint8)
        gts = np.zeros((n, H, W), dtype=bool)
        for i, f in enumerate(self.images_filenames):
            imgs[i] = load_image(str(f), self.imsize)
            gts[i] = get_ground_truth(get_ground_truth_filename(str(f), dataset_dir + "ground_truth/"), self.imsize)
        self.images = torch.from_numpy(imgs).to(self.device)
        self.gt = torch.from_numpy(gts).to(self.device)
        mean = torch.as_tensor(CPP.normalize_mean, dtype=torch.float32, device=self.device)[None, :, None, None]
        std = torch.as_tensor(CPP.normalize_std, dtype=torch.float32, device=self.device)[None, :, None, None]
        self._mean, self._std = mean, std

    def __len__(self) -> int:
        return int(self.images_filenames.shape[0])

    def get_batch(self, indices: Sequence[int]):
        idx = torch.as_tensor(np.asarray(indices), device=self.device)
        src = self.images.index_select(0, idx)
        original = to_tensor(src)                                    # ToTensor(original), datasets.py:81
        x = (original - self._mean) / self._std if self.transform else src   # Normalize: the same float32 sub / div
        gt = self.gt.index_select(0, idx)[:, None].to(torch.float32)         # ToTensor(gt) of a mode '1' image: 0.0 / 1.0
        return x, gt, original

    def __getitem__(self, index: int):
        x, gt, original = self.get_batch([int(index)])
        return x[0], gt[0], original[0]


class _EvalLoader:
    def __init__(self, dataset: MVTecDataset, batch_size: int, shuffle: bool, seed: int):
        self.dataset, self.batch_size, self.shuffle = dataset, int(batch_size), shuffle
        self.rng = np.random.default_rng(seed)

    def __len__(self) -> int:
        return (len(self.dataset) + self.batch_size - 1) // self.batch_size

    def __iter__(self) -> Iterator:
        n = len(self.dataset)
        order = self.rng.permutation(n) if self.shuffle else np.arange(n)
        for b in range(len(self)):
            yield self.dataset.get_batch(order[b * self.batch_size:(b + 1) * self.batch_size])


class MVTecDatamodule:
    """datasets.py:87-163 with the same arguments (``tools.inference`` builds it at tools.py:331-339)."""

    def __init__(self, root_dir: str, imsize: tuple = (256, 256), batch_size: int = 32, seed: int = 0, patch_level=None,
                 *, device="cuda"):
        self.root_dir, self.imsize, self.batch_size, self.seed, self.patch_level = root_dir, tuple(imsize), batch_size, seed, patch_level
        self.device = device
        self.train_images_filenames = get_filenames(self.root_dir + "/train/good/")
        self.test_images_filenames = get_test_data_filenames(self.root_dir + "/test/")

    def setup(self, stage=None) -> None:
        from sklearn.model_selection import train_test_split
        train, val = train_test_split(self.train_images_filenames, test_size=0.2, random_state=self.seed)
        mk = lambda names: MVTecDataset(self.root_dir, names, imsize=self.imsize, transform=True, device=self.device)   # noqa: E731
        self.train_dataset, self.val_dataset, self.test_dataset = mk(train), mk(val), mk(self.test_images_filenames)

    def train_dataloader(self):
        return _EvalLoader(self.train_dataset, self.batch_size, True, self.seed)

    def val_dataloader(self):
        return _EvalLoader(self.val_dataset, self.batch_size, False, self.seed)

    def test_dataloader(self):
        return _EvalLoader(self.test_dataset, self.batch_size, False, self.seed)

    def predict_dataloader(self):
        return _EvalLoader(self.test_dataset, self.batch_size, False, self.seed)
