"""-m gpu: ssad_obj_mask (canny -> morphology -> fill -> erosion -> largest component on device) against the CPU
restatement oracle/objmask.py (dataset_generator.py:27-39).  Bar: masks bit-exact."""
import numpy as np
import pytest
import torch

from ssad_b200 import synth
from ssad_b200.generator import obj_mask_batch, unpack_mask_bits

pytestmark = pytest.mark.gpu


def _run(images: np.ndarray) -> np.ndarray:
    src = torch.from_numpy(np.ascontiguousarray(images)).cuda()
    bits = obj_mask_batch(src)
    torch.cuda.synchronize()
    return unpack_mask_bits(bits.cpu().numpy(), images.shape[2])


def _oracle(images: np.ndarray) -> np.ndarray:
    from oracle import objmask
    return np.stack([objmask.obj_mask_array(im) for im in images])


def _check(images, what):
    got, want = _run(images), _oracle(images)
    for i in range(len(images)):
        diff = int(np.count_nonzero(got[i] != want[i]))
        assert diff == 0, f"{what}: image {i}: {diff} mask pixels differ (oracle area {int(want[i].sum())}, device {int(got[i].sum())})"


@pytest.mark.parametrize("size", [(128, 128), (256, 256), (96, 160), (100, 77), (64, 1024)])
def test_screw_like(size):
    H, W = size
    _check(synth.screw_images(6, H, W, 31 + H + W), f"screw {H}x{W}")


def test_object_and_texture_images():
    _check(synth.object_images(4, 192, 192, 5), "object")
    _check(synth.texture_images(4, 160, 224, 6), "texture")          # many edge fragments / components


def test_noise_many_components():
    rng = np.random.default_rng(3)
    img = rng.integers(0, 256, (3, 120, 150, 3), dtype=np.uint8)
    _check(img, "iid noise")
    blocks = np.kron(rng.integers(0, 2, (3, 12, 15, 1), dtype=np.uint8), np.ones((1, 10, 10, 3), np.uint8)) * 200 + 20
    _check(blocks.astype(np.uint8), "checker blocks")


def test_flat_image_degenerate_all_true():
    img = np.full((2, 64, 96, 3), 128, dtype=np.uint8)
    got = _run(img)
    assert got.all()                                                   # no edges -> label 0 wins -> all True
    _check(img, "flat")


def test_spiral_needs_many_sweeps():
    # a long spiral corridor: reachability has to reverse direction many times
    H = W = 128
    img = np.full((1, H, W, 3), 200, dtype=np.uint8)
    yy, xx = np.mgrid[0:H, 0:W]
    r = np.hypot(yy - 64, xx - 64)
    th = np.arctan2(yy - 64, xx - 64)
    spiral = (np.abs(((r - 6 * th / np.pi) % 12) - 6) < 1.5) & (r < 60)
    img[0][spiral] = 30
    _check(img, "spiral")


def test_1024_single():
    _check(synth.screw_images(1, 1024, 1024, 77), "screw 1024")


@pytest.mark.parametrize("knobs", [(30.0, -1, -1), (1.0, 0, -1), (8.0, 4, -1), (30.0, -1, 1), (300.0, -1, -1)],
                         ids=["wide_margins", "tiles_only", "small_tile_lists", "list_overflow", "everything_undecided"])
def test_float64_fallback_paths_give_the_same_masks(knobs):
    """ssad_obj_mask decides in float32 where its error margins allow, per pixel or per tile in float64 elsewhere: the
    masks must not depend on where a pixel was decided.  The knobs push pixels onto every path."""
    from ssad_b200 import _native
    lib = _native.load()
    try:
        lib.ssad_obj_mask_test_knobs(*knobs)
        _check(synth.screw_images(3, 192, 256, 91), "screw 192x256 " + str(knobs))
        _check(synth.texture_images(2, 160, 224, 7), "texture " + str(knobs))
        rng = np.random.default_rng(5)
        blocks = np.kron(rng.integers(0, 2, (2, 12, 15, 1), dtype=np.uint8), np.ones((1, 10, 10, 3), np.uint8)) * 200 + 20
        _check(blocks.astype(np.uint8), "checker blocks " + str(knobs))
    finally:
        lib.ssad_obj_mask_test_knobs(-1.0, -1, -1)


def _blob_images():
    """Bright blobs on a flat ground: several components of EQUAL size (the first in raster order must win), nested
    shapes, a comb whose teeth are one component, blobs one and two pixels apart, a blob on the image border."""
    H, W = 200, 288
    img = np.full((4, H, W, 3), 30, dtype=np.uint8)
    for (y, x) in [(20, 200), (20, 40), (120, 40), (120, 200)]:       # four equal squares
        img[0, y:y + 40, x:x + 40] = 220
    img[1, 30:170, 30:250] = 220                                       # ring with an island inside
    img[1, 60:140, 60:220] = 30
    img[1, 90:110, 100:180] = 220
    for k in range(12):                                                # comb: teeth joined at the bottom
        img[2, 20:150, 20 + 20 * k:30 + 20 * k] = 220
    img[2, 150:170, 20:250] = 220
    img[3, 0:50, 0:60] = 220                                           # on the border
    img[3, 80:130, 100:150] = 220                                      # two blobs a few pixels apart (closing may join them)
    img[3, 80:130, 153:203] = 220
    img[3, 150:190, 100:150] = 220
    return img


@pytest.mark.parametrize("max_runs", [-1, 0, 16])
def test_largest_component_paths_give_the_same_masks(max_runs):
    """The largest component is a union-find over the runs of the eroded plane when it holds at most `max_runs` runs, a
    peel of flooded components otherwise (include/ssad.h: ssad_obj_mask_test_runs): same masks on either path."""
    from ssad_b200 import _native
    lib = _native.load()
    try:
        lib.ssad_obj_mask_test_runs(max_runs)
        _check(_blob_images(), f"blobs max_runs={max_runs}")
        _check(synth.screw_images(3, 192, 256, 91), f"screw 192x256 max_runs={max_runs}")
        _check(synth.texture_images(2, 160, 224, 7), f"texture max_runs={max_runs}")
        rng = np.random.default_rng(5)
        blocks = np.kron(rng.integers(0, 2, (2, 12, 15, 1), dtype=np.uint8), np.ones((1, 10, 10, 3), np.uint8)) * 200 + 20
        _check(blocks.astype(np.uint8), f"checker blocks max_runs={max_runs}")
        _check(synth.screw_images(1, 1024, 1024, 77), f"screw 1024 max_runs={max_runs}")
    finally:
        lib.ssad_obj_mask_test_runs(-1)


def test_1024_batch_of_three():
    _check(synth.screw_images(3, 1024, 1024, 78), "screw 1024 x3")
