import numpy as np


def bits(a):
    return np.ascontiguousarray(a, dtype=np.float32).view(np.uint32)


def assert_bits_equal(a, b, what="", any_nan=False):
    """Bit-for-bit equality of two f32 arrays (NaN payloads included, unless any_nan: x86 and the
    GPU produce different default NaN encodings for 0/0, which no shader spec pins)."""
    ua, ub = bits(a), bits(b)
    assert ua.shape == ub.shape, (what, ua.shape, ub.shape)
    bad = ua != ub
    if any_nan:
        bad &= ~(np.isnan(np.asarray(a, np.float32)) & np.isnan(np.asarray(b, np.float32)))
    if bad.any():
        idx = np.argwhere(bad)
        first = tuple(idx[0])
        af, bf = np.asarray(a, np.float32), np.asarray(b, np.float32)
        raise AssertionError("%s: %d of %d f32 words differ; first at %s: %r vs %r" %
                             (what, int(bad.sum()), ua.size, first, af[first], bf[first]))


def scene_arrays(scene):
    cam, spheres, vertices, indices, meshes = scene
    return spheres, vertices, indices, meshes
