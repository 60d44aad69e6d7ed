"""CPU oracle for the artificial-anomaly generator hot path.

TEST INFRASTRUCTURE ONLY.  Nothing in the product package may import, call or
link anything under ``oracle/``; only ``tests/``, ``__graft_entry__.smoke()`` and
``bench.py``'s ``cpu_baseline`` / ``--impl reference`` legs do, and only as the
checker / the timed CPU baseline.

Contents
--------
``pillow_model``   numpy restatement of the third-party arithmetic the reference's
                   path executes inside Pillow 12.2.0 / torchvision 0.26.0 (paste,
                   16.16 nearest affine, polygon scan-line fill, blend, RGB->L);
                   pinned bit-exact against the installed libraries by
                   ``tests/test_pillow_model.py``.
``generator``      explicit-parameter restatement of the reference's per-image
                   pipeline (``src/self_supervised/datasets.py:209-394``,
                   ``src/self_supervised/dataset_generator.py``,
                   ``src/test_artificial_transformations.py:20-122``) driving the
                   installed Pillow/torchvision exactly as the reference does.
``objmask``        numpy/SciPy restatement of ``obj_mask``
                   (``dataset_generator.py:27-39``) including skimage's canny.
                   skimage is absent from this image and unpinned by the
                   reference => *parity unpinned* for this one function.
``reference_replay`` (only usable where ``/root/reference`` exists) runs the
                   reference's OWN code under a scripted RNG so that the same
                   explicit parameters produce the reference's own output; used by
                   ``tests/golden/make_golden.py`` to create the committed golden
                   vectors that pin ``generator`` (and the CUDA path).

Parity status: the reference ships no golden vectors or assertions for this path
(SURVEY.md section 4); the pin is (a) golden vectors generated from the reference's own
code in the build container (``tests/golden``), (b) the bit-exact checks of
``pillow_model`` against the installed Pillow.  ``objmask``: parity unpinned.
"""
