"""CPU-only checks of the C ABI: the library loads, exports every function include/ssad.h declares, the ctypes
binding covers the same set, and the struct layouts agree.  No kernel is launched."""
import ctypes as C
import os
import re

import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def declared_functions():
    text = open(os.path.join(ROOT, "include", "ssad.h")).read()
    text = re.sub(r"/\*.*?\*/", "", text, flags=re.S)
    return sorted(set(re.findall(r"\b(ssad_[a-z0-9_]+)\s*\(", text)))


def test_header_declares_the_expected_entry_points():
    names = declared_functions()
    for must in ("ssad_generate", "ssad_generate_workspace", "ssad_obj_mask", "ssad_polygon_mask", "ssad_affine_nearest",
                 "ssad_color_jitter", "ssad_select_kth", "ssad_paste", "ssad_to_tensor", "ssad_strerror", "ssad_version"):
        assert must in names


def test_library_exports_every_declared_symbol():
    import __graft_entry__ as g
    g.build()
    from ssad_b200 import _native
    lib = C.CDLL(_native.LIB_PATH)
    for name in declared_functions():
        assert hasattr(lib, name), f"{name} declared in include/ssad.h but not exported"
    assert sorted(_native.EXPORTS) == declared_functions()


def test_binding_loads_and_layouts_agree():
    from ssad_b200 import _native
    from ssad_b200.params import IMAGE_PARAMS_DTYPE
    lib = _native.load()
    assert lib.ssad_version() == 100
    assert lib.ssad_sizeof_params() == IMAGE_PARAMS_DTYPE.itemsize == 192
    assert lib.ssad_sizeof_desc() == C.sizeof(_native.SsadBatchDesc)
    assert lib.ssad_strerror(-1).decode() and lib.ssad_strerror(0).decode()


def test_argument_validation_happens_on_the_host():
    """Bad arguments are rejected before any CUDA call (so this runs without a GPU)."""
    from ssad_b200 import _native
    lib = _native.load()
    d = _native.SsadBatchDesc()
    assert lib.ssad_generate(None, *([None] * 10), 0, None) == -1
    d.batch, d.height, d.width, d.src_pitch = 1, 8, 8, 8          # pitch < 3 * width
    assert lib.ssad_generate(C.byref(d), C.c_void_p(16), C.c_void_p(16), *([None] * 8), 0, None) == -1
    assert lib.ssad_generate_workspace(None) == 0
    assert lib.ssad_generate_launches(None, 0) == 0
    with pytest.raises(ValueError):
        _native.check(-1, "x")
    with pytest.raises(RuntimeError):
        _native.check(-3, "x")
