"""The explicit-parameter oracle (``oracle/generator.py``) must reproduce, byte for
byte, the golden vectors the reference's own code produced
(``tests/golden/make_golden.py``)."""
import os

import numpy as np
import pytest

import scenarios as S
from golden import make_golden as G  # noqa: F401  (namespace package import of the generator script)

GOLD = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden")


def load_golden(name):
    g = np.load(os.path.join(GOLD, f"{name}.npz"))
    sc = S.build(name, G.BATCH[name])
    assert bytes(g["sha"]).hex() == G.scenario_sha(sc), \
        "synthetic inputs differ from the ones the golden vectors were made from: re-run tests/golden/make_golden.py"
    return sc, g


@pytest.mark.parametrize("name", S.ALL)
def test_oracle_reproduces_reference_golden(name):
    sc, g = load_golden(name)
    from oracle import generator as og
    from PIL import Image
    for i in range(len(sc.params)):
        x, y, _ = S.run_oracle(sc, i)
        assert y == int(g["y"][i])
        assert np.array_equal(x, g["x"][i]), (name, i)
        if i < 2:
            xn = og.to_tensor_normalize(Image.fromarray(x)).numpy()
            assert np.array_equal(xn, g["xn"][i])


def test_oracle_reproduces_script_function_golden():
    from PIL import Image
    from oracle import generator as og
    g = np.load(os.path.join(GOLD, "script_fns.npz"))
    sc = S.build("image_level", 8, seed=7)
    assert bytes(g["sha"]).hex() == G.scenario_sha(sc)
    seg = S.mask_to_pil(sc.mask)
    for which in ("do_patch", "do_scar"):
        for mode in ("image", "patch"):
            key = f"{which}_{mode}"
            draws = {k[len(key) + 3:]: g[k] for k in g.files if k.startswith(key + "_d_")}
            for i in range(len(sc.images)):
                d = {k: (v[i].copy() if v[i].ndim else v[i].item()) for k, v in draws.items()}
                img = Image.fromarray(sc.images[i])
                fn = og.do_patch if which == "do_patch" else og.do_scar
                x = fn(img, img, seg, mode == "patch", d)
                assert np.array_equal(np.array(x), g[key + "_x"][i]), (key, i)
