"""Class lists that drive the mask / warp policy of the generator.

Mirrors /root/reference/src/self_supervised/constants.py:84-119 (only the lists the
hot path reads: ``TEXTURES``, ``OBJECTS``, ``NON_FIXED_OBJECTS``).
"""


def TEXTURES() -> list:
    return ['carpet', 'grid', 'leather', 'tile', 'wood']


def OBJECTS() -> list:
    return ['bottle', 'cable', 'capsule', 'hazelnut', 'metal_nut', 'pill', 'screw', 'tile',
            'toothbrush', 'transistor', 'zipper']


def NON_FIXED_OBJECTS() -> list:
    return ['hazelnut', 'screw', 'metal_nut']


# class-specific pre-crop applied in patch-level mode (datasets.py:243-248), PIL box (l, t, r, b)
PATCH_MODE_PRECROP = {
    'capsule': (0, 50, 255, 200),
    'screw': (25, 25, 230, 230),
}
