"""Image comparison of SURVEY §8(c)/(f3) (host/rt_imgdiff.cpp): RMSE and mean LDR-FLIP.

The reference has no comparison tool; NVIDIA's FLIP implementation is not in this image.  The C++
restatement is checked against an independent numpy/scipy restatement of the paper's equations written
here, and against closed-form cases."""
import numpy as np
import pytest
from scipy import ndimage

from ray_tracer_b200 import host

PPD = host.FLIP_PPD


def _lin(s):
    return np.where(s > 0.04045, ((s + 0.055) / 1.055) ** 2.4, s / 12.92)


M = np.array([[10135552 / 24577794, 8788810 / 24577794, 4435075 / 24577794],
              [2613072 / 12288897, 8788810 / 12288897, 887015 / 12288897],
              [1425312 / 73733382, 8788810 / 73733382, 70074185 / 73733382]])
MI = np.array([[3.241003275, -1.537398934, -0.498615861], [-0.969224334, 1.875930071, 0.041554224],
               [0.055639423, -0.204011202, 1.057148933]])
WHITE = M @ np.ones(3)


def _ycxcz(xyz):
    n = xyz / WHITE
    return np.stack([116 * n[..., 1] - 16, 500 * (n[..., 0] - n[..., 1]), 200 * (n[..., 1] - n[..., 2])], -1)


def _lab(xyz):
    n = xyz / WHITE
    d = 6 / 29
    f = np.where(n > 0.00885, np.cbrt(n), n / (3 * d * d) + 4 / 29)
    return np.stack([116 * f[..., 1] - 16, 500 * (f[..., 0] - f[..., 1]), 200 * (f[..., 1] - f[..., 2])], -1)


def _hunt(lab):
    return np.stack([lab[..., 0], 0.01 * lab[..., 0] * lab[..., 1], 0.01 * lab[..., 0] * lab[..., 2]], -1)


def _hyab(a, b):
    return np.abs(a[..., 0] - b[..., 0]) + np.hypot(a[..., 1] - b[..., 1], a[..., 2] - b[..., 2])


def _csf(a1, b1, a2, b2, r):
    x = np.arange(-r, r + 1) / PPD
    d = x[None, :] ** 2 + x[:, None] ** 2
    g = a1 * np.sqrt(np.pi / b1) * np.exp(-np.pi ** 2 * d / b1) + a2 * np.sqrt(np.pi / b2) * np.exp(-np.pi ** 2 * d / b2)
    return g / g.sum()


def flip_numpy(ref, test):
    """LDR-FLIP from the paper's equations: per-pixel error map."""
    yr, yt = _ycxcz(_lin(ref) @ M.T), _ycxcz(_lin(test) @ M.T)
    r = int(np.ceil(3 * np.sqrt(0.04 / (2 * np.pi ** 2)) * PPD))
    ks = [_csf(1, 0.0047, 0, 1e-5, r), _csf(1, 0.0053, 0, 1e-5, r), _csf(34.1, 0.04, 13.5, 0.025, r)]

    def perceptual(y):
        f = np.stack([ndimage.correlate(y[..., c], ks[c], mode="nearest") for c in range(3)], -1)
        yy = (f[..., 0] + 16) / 116
        xyz = np.stack([yy + f[..., 1] / 500, yy, yy - f[..., 2] / 200], -1) * WHITE
        return _hunt(_lab(np.clip(xyz @ MI.T, 0, 1) @ M.T))
    qc, pc, pt = 0.7, 0.4, 0.95
    cmax = _hyab(_hunt(_lab(M @ np.array([0, 1.0, 0]))), _hunt(_lab(M @ np.array([0, 0, 1.0])))) ** qc
    d = _hyab(perceptual(yr), perceptual(yt)) ** qc
    dec = np.where(d < pc * cmax, pt / (pc * cmax) * d, pt + (d - pc * cmax) / (cmax - pc * cmax) * (1 - pt))
    sd = 0.5 * 0.082 * PPD
    rf = int(np.ceil(3 * sd))
    gx, gy = np.meshgrid(np.arange(-rf, rf + 1), np.arange(-rf, rf + 1))
    g = np.exp(-(gx ** 2 + gy ** 2) / (2 * sd * sd))

    def norm(k):
        return np.where(k > 0, k / k[k > 0].sum(), k / -k[k < 0].sum())
    edge, point = norm(-gx * g), norm((gx ** 2 / sd ** 2 - 1) * g)

    def mag(y, k):
        a = (y[..., 0] + 16) / 116
        return np.hypot(ndimage.correlate(a, k, mode="nearest"), ndimage.correlate(a, k.T, mode="nearest"))
    de = np.abs(mag(yr, edge) - mag(yt, edge))
    dp = np.abs(mag(yr, point) - mag(yt, point))
    def_ = (np.maximum(de, dp) / np.sqrt(2)) ** 0.5
    return dec ** (1 - def_)


def _pair(seed, h=48, w=64):
    rng = np.random.default_rng(seed)
    yy, xx = np.mgrid[0:h, 0:w]
    base = np.stack([0.5 + 0.4 * np.sin(xx / 7.0), 0.5 + 0.4 * np.cos(yy / 5.0), (xx + yy) / (h + w)], -1)
    base[h // 3: 2 * h // 3, w // 4: w // 2] = (0.9, 0.2, 0.1)            # an edge
    ref = np.clip(base, 0, 1).astype(np.float32)
    test = np.clip(base + rng.normal(0, 0.05, base.shape), 0, 1).astype(np.float32)
    test[5, 9] = (1, 1, 1)                                              # a firefly
    return ref, test


def test_flip_matches_numpy_restatement():
    for seed in (1, 2):
        ref, test = _pair(seed)
        mean, emap = host.image_flip(ref, test, want_map=True)
        want = flip_numpy(ref.astype(np.float64), test.astype(np.float64))
        assert emap.shape == want.shape
        assert np.max(np.abs(emap - want)) < 2e-5
        assert abs(mean - want.mean()) < 1e-6
        assert 0.0 < mean < 1.0


def test_flip_closed_form_cases():
    ref, test = _pair(3)
    assert host.image_flip(ref, ref) == 0.0
    assert host.image_flip(test, test) == 0.0
    a, b = host.image_flip(ref, test), host.image_flip(test, ref)
    assert abs(a - b) < 1e-12                                           # every term is a symmetric difference
    # uniform black vs uniform white: no features, dEc alone.  HyAB = 100, 100^0.7 = 25.12 lies above pc*cmax,
    # so the map is pt + (25.12 - 0.4 cmax) / (0.6 cmax) * 0.05 with cmax = HyAB(green, blue)^0.7
    black, white = np.zeros((12, 16, 3), np.float32), np.ones((12, 16, 3), np.float32)
    cmax = _hyab(_hunt(_lab(M @ np.array([0, 1.0, 0]))), _hunt(_lab(M @ np.array([0, 0, 1.0])))) ** 0.7
    want = 0.95 + (100 ** 0.7 - 0.4 * cmax) / (0.6 * cmax) * 0.05
    assert abs(host.image_flip(black, white) - want) < 1e-7        # the 9-digit inverse matrix of the metric is not an exact inverse
    assert 0.96 < want < 0.97
    # uint8 input is the same image scaled by 1/255
    u = (ref * 255).round().astype(np.uint8)
    v = (test * 255).round().astype(np.uint8)
    assert abs(host.image_flip(u, v) - host.image_flip(u.astype(np.float32) / np.float32(255), v.astype(np.float32) / np.float32(255))) < 1e-12


def test_rmse_and_tonemap():
    rng = np.random.default_rng(0)
    a = rng.random((9, 7, 4)).astype(np.float32)
    b = rng.random((9, 7, 4)).astype(np.float32)
    assert abs(host.image_rmse(a, b) - np.sqrt(np.mean((a.astype(np.float64) - b) ** 2))) < 1e-12
    assert host.image_rmse(a, a) == 0.0
    with pytest.raises(host.HostError):
        host.image_rmse(a, b[:-1])
    with pytest.raises(host.HostError):
        host.image_flip(a, b[:-1])
    t = host.tonemap(np.array([[[-1.0, 0.0, 0.002, 1.0], [0.5, 1.0, 7.0, 1.0]]], np.float32))
    want = [[0.0, 0.0, 0.002 * 12.92], [1.055 * 0.5 ** (1 / 2.4) - 0.055, 1.0, 1.0]]
    assert np.allclose(t[0], want, atol=1e-7)


def test_imgdiff_cli(tmp_path):
    from ray_tracer_b200 import imgdiff
    ref, test = _pair(4)
    pa, pb, pm = str(tmp_path / "a.png"), str(tmp_path / "b.png"), str(tmp_path / "m.png")
    for p, im in ((pa, ref), (pb, test)):
        rgba = np.concatenate([(im * 255).round().astype(np.uint8), np.full(im.shape[:2] + (1,), 255, np.uint8)], -1)
        host.write_png(p, rgba)
    out = imgdiff.main([pa, pb, "--map", pm])
    u, v = host.load_png(pa), host.load_png(pb)
    assert abs(out["flip_mean"] - host.image_flip(u, v)) < 1e-12
    assert abs(out["rmse"] - host.image_rmse(u[..., :3].astype(np.float32) / 255, v[..., :3].astype(np.float32) / 255)) < 1e-6
    m = host.load_png(pm)
    assert m.shape == u.shape and m[5, 9, 0] > m[40, 60, 0]             # the firefly is the bright spot of the map
    assert imgdiff.main([pa, pa])["flip_mean"] == 0.0
