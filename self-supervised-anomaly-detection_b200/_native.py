"""ctypes binding of libssad_gen.so (the C ABI in include/ssad.h).

There is no CPU fallback: if the library is missing or a symbol is absent this module raises."""
from __future__ import annotations

import ctypes as C
import os

HERE = os.path.dirname(os.path.abspath(__file__))
LIB_PATH = os.environ.get("SSAD_LIB") or os.path.join(HERE, "lib", "libssad_gen.so")   # SSAD_LIB: experiments only

SSAD_MASK_ALL, SSAD_MASK_SHARED, SSAD_MASK_PER_IMAGE = 0, 1, 2
SSAD_NORM_NONE, SSAD_NORM_F32, SSAD_NORM_BF16 = 0, 1, 2


class SsadBatchDesc(C.Structure):
    _fields_ = [
        ("batch", C.c_int32), ("height", C.c_int32), ("width", C.c_int32),
        ("src_height", C.c_int32), ("src_width", C.c_int32),
        ("src_pitch", C.c_int64), ("src_stride", C.c_int64),
        ("donor_count", C.c_int32), ("donor_height", C.c_int32), ("donor_width", C.c_int32),
        ("donor_pitch", C.c_int64), ("donor_stride", C.c_int64),
        ("donor_win_h", C.c_int32), ("donor_win_w", C.c_int32),
        ("mask_mode", C.c_int32), ("mask_pitch_words", C.c_int32), ("mask_stride_words", C.c_int64),
        ("mask_height", C.c_int32), ("mask_width", C.c_int32),
        ("box_poly", C.c_int32 * 4), ("box_scar", C.c_int32 * 4),
        ("coverage_min3", C.c_int32), ("max_patch_w", C.c_int32), ("max_patch_h", C.c_int32),
        ("out_norm_dtype", C.c_int32), ("crop_align16", C.c_int32),
        ("out_u8_pitch", C.c_int64), ("out_u8_stride", C.c_int64),
        ("any_warp", C.c_int32), ("sat_factor_max", C.c_float),
        ("src_indexed", C.c_int32), ("src_count", C.c_int32),
        ("line_width", C.c_int32), ("line_max", C.c_int32), ("line_xy", C.c_void_p),
        ("max_scar_w", C.c_int32), ("max_scar_h", C.c_int32),
    ]


class SsadCannyConsts(C.Structure):
    _fields_ = [("gauss_w", C.c_double * 13), ("low", C.c_double), ("high", C.c_double), ("eps", C.c_double)]


EXPORTS = {
    # name: (restype, argtypes)
    "ssad_version": (C.c_int, []),
    "ssad_strerror": (C.c_char_p, [C.c_int]),
    "ssad_last_cuda_error": (C.c_char_p, []),
    "ssad_sizeof_params": (C.c_size_t, []),
    "ssad_sizeof_desc": (C.c_size_t, []),
    "ssad_sizeof_debug": (C.c_size_t, []),
    "ssad_generate": (C.c_int, [C.POINTER(SsadBatchDesc)] + [C.c_void_p] * 10 + [C.c_size_t, C.c_void_p]),
    "ssad_generate_ctx": (C.c_int, [C.c_void_p, C.POINTER(SsadBatchDesc)] + [C.c_void_p] * 10 + [C.c_size_t, C.c_void_p]),
    "ssad_context_create": (C.c_int, [C.POINTER(C.c_void_p)]),
    "ssad_context_destroy": (None, [C.c_void_p]),
    "ssad_reload_env": (None, []),
    "ssad_generate_workspace": (C.c_size_t, [C.POINTER(SsadBatchDesc)]),
    "ssad_generate_launches": (C.c_int, [C.POINTER(SsadBatchDesc), C.c_int]),
    "ssad_obj_mask": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                C.c_int32, C.c_int64, C.c_void_p, C.c_size_t, C.c_void_p]),
    "ssad_obj_mask_workspace": (C.c_size_t, [C.c_int32, C.c_int32, C.c_int32]),
    "ssad_obj_mask_launches": (C.c_int, []),
    "ssad_obj_mask_test_knobs": (None, [C.c_float, C.c_int32, C.c_int32]),
    "ssad_obj_mask_test_runs": (None, [C.c_int32]),
    "ssad_upload_windows": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int64, C.c_int64, C.c_int32, C.c_int32,
                                      C.c_int32, C.c_int32, C.c_int32, C.c_void_p]),
    "ssad_obj_mask_static": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p,
                                       C.c_int32, C.c_int64, C.c_void_p]),
    "ssad_polygon_mask": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "ssad_line_mask": (C.c_int, [C.c_int32, C.c_void_p, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_void_p]),
    "ssad_affine_nearest": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.POINTER(C.c_int32), C.c_void_p,
                                      C.c_int32, C.c_int32, C.c_void_p]),
    "ssad_color_jitter": (C.c_int, [C.c_int32, C.c_int32, C.c_int32] + [C.c_void_p] * 6),
    "ssad_select_kth": (C.c_int, [C.c_void_p] + [C.c_int32] * 6 + [C.c_void_p] * 4),
    "ssad_paste": (C.c_int, [C.c_void_p, C.c_int32, C.c_int32, C.c_void_p, C.c_int32, C.c_int32, C.c_int32, C.c_void_p,
                             C.c_int32, C.c_int32, C.c_void_p]),
    "ssad_color_similarity": (C.c_int, [C.c_void_p, C.c_int32, C.c_void_p, C.c_int32, C.c_void_p, C.c_void_p]),
    "ssad_to_tensor": (C.c_int, [C.c_int32, C.c_int32, C.c_int32, C.c_void_p, C.c_int64, C.c_int64, C.c_void_p, C.c_void_p]),
}

_lib = None


def load() -> C.CDLL:
    """Load the shared library and bind every symbol include/ssad.h declares."""
    global _lib
    if _lib is not None:
        return _lib
    if not os.path.exists(LIB_PATH):
        raise RuntimeError(
            f"{LIB_PATH} not found: build it with `python -c 'import __graft_entry__ as g; g.build()'` "
            "(nvcc, sm_100a). There is no CPU fallback.")
    lib = C.CDLL(LIB_PATH)
    for name, (res, args) in EXPORTS.items():
        fn = getattr(lib, name)      # AttributeError if the symbol is missing
        fn.restype = res
        fn.argtypes = args
    from .params import IMAGE_PARAMS_DTYPE
    if lib.ssad_sizeof_params() != IMAGE_PARAMS_DTYPE.itemsize:
        raise RuntimeError("SsadImageParams layout mismatch between include/ssad.h and params.py")
    if lib.ssad_sizeof_desc() != C.sizeof(SsadBatchDesc):
        raise RuntimeError("SsadBatchDesc layout mismatch between include/ssad.h and _native.py")
    _lib = lib
    return lib


def check(code: int, what: str) -> None:
    if code == 0:
        return
    lib = load()
    msg = lib.ssad_strerror(code).decode()
    if code == -4:
        msg += ": " + lib.ssad_last_cuda_error().decode()
        raise RuntimeError(f"{what}: {msg}")
    if code in (-1, -2):
        raise ValueError(f"{what}: {msg}")
    raise RuntimeError(f"{what}: {msg}")
