"""CPU: the vectorised host derivations of ``params.derive_records`` are byte-identical to the per-image scalar
routines (``random_affine_matrix`` / ``rotate_expand_matrix`` / ``polygon_vertices`` -- Python floats through math.*,
the way torchvision and PIL/Image.py compute them) for the same ``numpy.random.Generator`` state."""
import numpy as np
import pytest

import ssad_b200
from ssad_b200 import GeneratorConfig
from ssad_b200 import params as P

CONFIGS = [
    dict(subject="bottle"),
    dict(subject="bottle", imsize=(320, 320), patch_localization=True, patch_size=256,
         patch_area_ratio=(0.1, 0.2), scar_area_ratio=(0.01, 0.02)),
    dict(subject="carpet", n_donors=15),
    dict(subject="screw", imsize=(1024, 1024)),
    dict(subject="capsule", patch_localization=True, patch_size=32),
]


def _scalar_fields(cfg, draws, rec):
    """warp_fix / poly / scar fields of the records, one image at a time with the scalar routines."""
    B = len(rec)
    W, H = cfg.work_size
    out = {k: np.zeros_like(rec[k]) for k in ("warp_fix", "poly_n", "poly_xy", "scar_fix", "scar_size")}
    pw, ph = rec["rect"][:, 2], rec["rect"][:, 3]
    for i in range(B):
        if cfg.warp:
            m = P.random_affine_matrix(W, H, float(draws["warp_angle"][i]), float(draws["warp_scale"][i]))
            out["warp_fix"][i] = P.affine_fixed_coeffs(m)
        if draws["label"][i] == 1:
            pts = P.polygon_vertices(int(pw[i]), int(ph[i]), draws["poly_n_side"][i], draws["poly_v"][i])
            out["poly_n"][i] = len(pts)
            out["poly_xy"][i, :len(pts)] = pts
        if draws["label"][i] == 2:
            m, nw, nh = P.rotate_expand_matrix(int(pw[i]), int(ph[i]), float(draws["scar_angle"][i]))
            out["scar_fix"][i] = P.affine_fixed_coeffs(m)
            out["scar_size"][i] = (nw, nh)
    return out


@pytest.mark.parametrize("kw", CONFIGS, ids=lambda k: k["subject"] + ("_patch" if k.get("patch_localization") else ""))
def test_vectorised_records_equal_scalar_derivation(kw):
    cfg = GeneratorConfig(**kw)
    pb = P.sample_params(cfg, 600, np.random.default_rng(42))
    want = _scalar_fields(cfg, pb.draws, pb.records)
    for k, v in want.items():
        assert np.array_equal(pb.records[k], v), k


def test_rotation_covers_every_scar_angle_and_identity():
    w = np.array([7, 12, 39, 2, 156], dtype=np.int64)
    h = np.array([3, 30, 11, 2, 41], dtype=np.int64)
    for ang in list(range(-45, 46)) + [0, 90, 180, 270, 360, -360]:
        fix, size = P.rotate_expand_fixed(w, h, np.full(len(w), ang))
        for i in range(len(w)):
            m, nw, nh = P.rotate_expand_matrix(int(w[i]), int(h[i]), float(ang))
            assert tuple(fix[i]) == P.affine_fixed_coeffs(m), (ang, i)
            assert tuple(size[i]) == (nw, nh), (ang, i)


def test_sampler_has_no_per_image_python_loop():
    """The per-image Python loop of round 1 drew 52 k records/s at best; the vectorised derivations are several times
    faster at large batches (generous bound: the CPU test box is shared)."""
    import time
    cfg = GeneratorConfig(**CONFIGS[1])
    rng = np.random.default_rng(0)
    P.sample_params(cfg, 64, rng)
    best = 0.0
    for _ in range(5):
        t = time.perf_counter()
        P.sample_params(cfg, 4096, rng)
        best = max(best, 4096 / (time.perf_counter() - t))
    assert best > 100e3, f"{best:.0f} records/s"
