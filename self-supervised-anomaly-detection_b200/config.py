"""Generator configuration: the reference's ``CPP`` constants and the per-dataset
settings ``PretextTaskDataset.__init__`` derives (datasets.py:33-47,166-206)."""
from __future__ import annotations

from dataclasses import dataclass, field
from typing import Optional, Tuple

from . import constants


class CPP:
    """datasets.py:33-47.  ``jitter_transforms`` is a torchvision object in the
    reference; here it is the batched device op (see ``datasets.py`` of this package),
    the numeric constants are identical."""
    jitter_offset = 0.1

    rectangle_area_ratio_patch = (0.2, 0.5)   # patch-wise
    rectangle_area_ratio = (0.03, 0.07)       # image-wise
    rectangle_aspect_ratio = ((0.3, 0.5), (1, 3.3))

    scar_area_ratio_patch = (0.02, 0.05)      # patch-wise
    scar_area_ratio = (0.003, 0.007)          # image-wise
    scar_aspect_ratio = ((0.3, 0.5), (2.5, 3.3))

    # datasets.py:238-239
    container_scaling_factor_patch = 1.75
    container_scaling_factor_scar = 2
    # datasets.py:270,311
    variant_probabilities = (0.7, 0.15, 0.15)
    # datasets.py:295-297
    similarity_threshold = 0.99
    brightness_low = (0.75, 0.9)
    brightness_high = (1.1, 1.15)
    # datasets.py:221
    affine_degrees = 3
    affine_scale = (1.05, 1.1)
    # datasets.py:430-433
    normalize_mean = (0.485, 0.456, 0.406)
    normalize_std = (0.229, 0.224, 0.225)


@dataclass
class GeneratorConfig:
    """Everything ``__getitem__`` reads from ``self`` / constants.

    ``imsize`` is ``(W, H)`` as in the reference (PIL order).  In patch-level mode the
    working image is ``patch_size x patch_size``.
    """
    subject: str = 'bottle'
    imsize: Tuple[int, int] = (256, 256)
    patch_localization: bool = False
    patch_size: int = 64
    num_labels: int = 3                     # 3-class contract {0,1,2}; 4 = HEAD's randint(0,3) incl. label 3 "long line"
    jitter: bool = True
    n_donors: int = 0                       # len(images_for_cut) (textures only)
    patch_area_ratio: Optional[Tuple[float, float]] = None
    scar_area_ratio: Optional[Tuple[float, float]] = None
    rectangle_aspect_ratio: Tuple[Tuple[float, float], Tuple[float, float]] = CPP.rectangle_aspect_ratio
    scar_aspect_ratio: Tuple[Tuple[float, float], Tuple[float, float]] = CPP.scar_aspect_ratio
    # function-level entry points (do_patch / do_scar, test_artificial_transformations.py:20-122) run the image-level
    # geometry without the pre-warp and, with patch_loc=True, with container factor 1
    force_warp: Optional[bool] = None
    force_container_factor: Optional[float] = None
    force_line_patch_mode: Optional[bool] = None   # do_swirl(patch_loc=True): 30 points, width 1, no split (script :131,:152-164)

    def __post_init__(self):
        if not 1 <= int(self.num_labels) <= 4:
            raise ValueError("num_labels must be in 1..4 (datasets.py:215 draws randint(0, 3))")
        if self.patch_area_ratio is None:   # datasets.py:184-185
            self.patch_area_ratio = CPP.rectangle_area_ratio_patch if self.patch_localization else CPP.rectangle_area_ratio
        if self.scar_area_ratio is None:
            self.scar_area_ratio = CPP.scar_area_ratio_patch if self.patch_localization else CPP.scar_area_ratio
        if isinstance(self.patch_size, (tuple, list)):
            self.patch_size = int(self.patch_size[0])
        self.imsize = (int(self.imsize[0]), int(self.imsize[1]))

    # -- class policy (datasets.py:197,220,225,232)
    @property
    def is_texture(self) -> bool:
        return self.subject in constants.TEXTURES()

    @property
    def is_non_fixed(self) -> bool:
        return self.subject in constants.NON_FIXED_OBJECTS()

    @property
    def warp(self) -> bool:
        if self.force_warp is not None:
            return bool(self.force_warp)
        return (not self.patch_localization) and (not self.is_non_fixed)

    @property
    def line_patch_mode(self) -> bool:
        """Label 3: 30 picks / width 1 / no split (patch mode) or 60 picks / width 3 / one tenth (datasets.py:360,374-388)."""
        if self.force_line_patch_mode is not None:
            return bool(self.force_line_patch_mode)
        return bool(self.patch_localization)

    # -- geometry
    @property
    def precrop(self) -> Tuple[int, int, int, int]:
        """PIL box applied to x and the mask before the random window (patch mode)."""
        W, H = self.imsize
        if self.patch_localization and self.subject in constants.PATCH_MODE_PRECROP:
            return constants.PATCH_MODE_PRECROP[self.subject]
        return (0, 0, W, H)

    @property
    def work_size(self) -> Tuple[int, int]:
        """(W, H) of the working image x."""
        if self.patch_localization:
            return (self.patch_size, self.patch_size)
        return self.imsize

    @property
    def donor_window(self) -> Tuple[int, int]:
        """(W, H) of image_for_cutting as seen by generate_patch."""
        return self.work_size

    def container(self, factor: float) -> Tuple[int, int, int, int]:
        """dataset_generator.py:15-24 for x.size = work_size (uses size[0] only)."""
        c = int(self.work_size[0] / 2)
        lo = int(c - (c / factor))
        hi = int(c + (c / factor))
        return (lo, lo, hi, hi)

    @property
    def box_poly(self):
        if self.force_container_factor is not None:
            return self.container(self.force_container_factor)
        return self.container(1 if self.patch_localization else CPP.container_scaling_factor_patch)

    @property
    def box_scar(self):
        if self.force_container_factor is not None:
            return self.container(self.force_container_factor)
        return self.container(1 if self.patch_localization else CPP.container_scaling_factor_scar)

    @property
    def coverage_min3(self) -> int:
        """datasets.py:258: y=0 iff sum(ToTensor(seg_rgb)) = 3*count < int(ps*ps/2)."""
        if not self.patch_localization:
            return 0
        return int((self.patch_size * self.patch_size) / 2)
