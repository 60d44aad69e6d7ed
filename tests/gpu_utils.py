"""Helpers for the -m gpu parity tests: run a Scenario through the C-ABI CUDA path."""
import numpy as np
import torch

import scenarios as S
from ssad_b200.generator import BatchGenerator, pack_mask_bits


def make_generator(sc: S.Scenario) -> BatchGenerator:
    fixed = sc.mask if (sc.mask is not None and sc.mask.ndim == 2) else None
    return BatchGenerator(sc.cfg, "cuda", mask=fixed, donors=sc.donors)


def per_image_masks(sc: S.Scenario):
    if sc.mask is not None and sc.mask.ndim == 3:
        return torch.from_numpy(pack_mask_bits(sc.mask).view(np.int32)).cuda()
    return None


def run_cuda(sc: S.Scenario, norm_dtype=torch.float32, debug=True, gen=None):
    gen = gen or make_generator(sc)
    src = torch.from_numpy(sc.images).cuda()
    out = gen.generate(src, sc.params, masks=per_image_masks(sc), want_u8=True, norm_dtype=norm_dtype, debug=debug)
    torch.cuda.synchronize()
    return out
