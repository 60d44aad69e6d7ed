"""-m gpu: ``ssad_obj_mask_static`` (include/ssad.h) -- ssad_obj_mask with its scratch in the library's static device
array, for callers that cannot allocate -- gives the masks of ``ssad_obj_mask`` bit for bit, on calls in a row that share
the array, and rejects requests beyond SSAD_OBJ_MASK_STATIC_BYTES.  Called through the C ABI (ctypes); the only torch
allocations are the inputs and outputs, made before the calls.  (Collected last on purpose: pytest -x.)"""
import ctypes as C

import numpy as np
import pytest
import torch

from ssad_b200 import _native, synth
from ssad_b200.generator import canny_consts, obj_mask_batch

pytestmark = pytest.mark.gpu

STATIC_BYTES = 16 << 20          # SSAD_OBJ_MASK_STATIC_BYTES


def _static(lib, src, out):
    B, H, W = (int(v) for v in src.shape[:3])
    k = canny_consts()
    with torch.cuda.device(src.device):
        stream = torch.cuda.current_stream(src.device).cuda_stream
        return lib.ssad_obj_mask_static(B, H, W, C.c_void_p(src.data_ptr()), int(src.stride(1)), int(src.stride(0)), C.byref(k),
                                        C.c_void_p(out.data_ptr()), int(out.stride(1)), int(out.stride(0)), C.c_void_p(stream))


def test_static_scratch_entry_point_matches_obj_mask_on_calls_in_a_row():
    lib = _native.load()
    batches = [synth.screw_images(5, 192, 256, 91),                 # three calls of different shapes share the array:
               synth.texture_images(3, 160, 224, 7),                # the second and third find the previous call's planes,
               synth.screw_images(2, 1024, 1024, 77),               # lists and counters in it
               synth.screw_images(5, 192, 256, 91)]
    srcs = [torch.from_numpy(np.ascontiguousarray(b)).cuda() for b in batches]
    want = [obj_mask_batch(s) for s in srcs]                        # ssad_obj_mask with a torch workspace
    outs = [torch.full_like(w, -1) for w in want]
    torch.cuda.synchronize()
    for rep in range(2):
        for s, o in zip(srcs, outs):
            assert int(lib.ssad_obj_mask_workspace(*[int(v) for v in s.shape[:3]])) <= STATIC_BYTES
            before = torch.cuda.memory_allocated()
            rc = _static(lib, s, o)
            assert rc == 0, lib.ssad_strerror(rc).decode()
            assert torch.cuda.memory_allocated() == before
        torch.cuda.synchronize()
        for i, (o, w) in enumerate(zip(outs, want)):
            assert torch.equal(o, w), f"pass {rep}, batch {i}: {int((o != w).sum())} mask words differ"
            o.fill_(-1)
    assert torch.equal(outs[0], torch.full_like(outs[0], -1))


def test_static_scratch_entry_point_rejects_what_does_not_fit():
    lib = _native.load()
    src = torch.zeros((1, 64, 64, 3), dtype=torch.uint8, device="cuda")
    out = torch.full((1, 64, 2), 7, dtype=torch.int32, device="cuda")
    k = canny_consts()
    B, H, W = 64, 1024, 1024                                        # 64 x 425 KB of planes and lists > 16 MiB
    assert int(lib.ssad_obj_mask_workspace(B, H, W)) > STATIC_BYTES
    rc = lib.ssad_obj_mask_static(B, H, W, C.c_void_p(src.data_ptr()), 3 * W, 3 * W * H, C.byref(k), C.c_void_p(out.data_ptr()),
                                  32, 32 * H, C.c_void_p(0))
    assert rc == -6                                                 # SSAD_ESTATIC, nothing launched
    assert "static" in lib.ssad_strerror(rc).decode()
    torch.cuda.synchronize()
    assert bool((out == 7).all())
    with pytest.raises(RuntimeError, match="static"):
        _native.check(rc, "ssad_obj_mask_static")
