"""Texture extension (no reference counterpart — upstream has no UVs and loads no images):
oracle-side pins of the sampling rule and of "zero padding == reference behaviour", and GPU
parity of the CUDA texture-object path against the oracle."""
import numpy as np
import pytest

from util import assert_bits_equal


def _checker():
    t = np.zeros((2, 2, 4), np.uint8)
    t[0, 0], t[0, 1], t[1, 0], t[1, 1] = (255, 0, 0, 255), (0, 255, 0, 255), (0, 0, 255, 255), (255, 255, 255, 255)
    return t


def test_sampling_rule(oracle):
    t = _checker()
    k = np.float32(0.003921568859368563)
    # texel centres return the texel, channel = byte * f32(1/255)
    assert np.array_equal(oracle.sample_texture(t, 0.25, 0.25), np.float32([255, 0, 0, 255]) * k)
    assert np.array_equal(oracle.sample_texture(t, 0.75, 0.75), np.float32([255, 255, 255, 255]) * k)
    # half way between two centres: mix(a, b, 0.5)
    assert np.allclose(oracle.sample_texture(t, 0.5, 0.25), [0.5, 0.5, 0, 1], atol=1e-7)
    # repeat addressing: (u, v) and (u + k, v - m) agree; the corner blends all four texels
    assert np.array_equal(oracle.sample_texture(t, 0.25, 0.25), oracle.sample_texture(t, -0.75, 1.25))
    assert np.allclose(oracle.sample_texture(t, 0.0, 0.0), [0.5, 0.5, 0.5, 1.0], atol=1e-7)
    # non-finite coordinates fall back to texel 0 instead of indexing out of range
    assert np.isfinite(oracle.sample_texture(t, np.inf, np.nan)).all()
    # non-square, non-power-of-two
    r = np.random.default_rng(3).integers(0, 256, (5, 7, 4), dtype=np.uint8)
    for (i, j) in ((0, 0), (6, 4), (3, 2)):
        assert np.array_equal(oracle.sample_texture(r, (i + 0.5) / 7, (j + 0.5) / 5), r[j, i].astype(np.float32) * k)


def _cfg(name, w, h):
    from ray_tracer_b200 import configs
    from oracle import rt_oracle as ro
    scene, meta = configs.build(name, w, h)
    t = scene.as_tuple()
    return scene, (t[0].view(ro.CAMERA_DT), t[1].view(ro.SPHERE_DT), t[2].view(ro.VERTEX_DT), t[3], t[4].view(ro.MESH_DT))


def test_zero_handles_are_reference_behaviour(oracle):
    """UVs without handles, and handles without a texture table, change nothing."""
    scene, sc = _cfg("c2_cube", 40, 24)
    _, plain = _cfg("c2_cube_plain", 40, 24)
    base = oracle.render(plain, 40, 24, 1, 4, 1, 1)
    cam, sp, vs, ix, ms = sc
    ms0 = ms.copy(); ms0["diffuse_tex"] = 0; ms0["normal_tex"] = 0; ms0["color"] = plain[4]["color"]; ms0["specular"] = plain[4]["specular"]
    with_uv = oracle.render((cam, sp, vs, ix, ms0), 40, 24, 1, 4, 1, 1, textures=scene.textures)
    assert_bits_equal(with_uv, base, "uvs + table but no handles")
    ms1 = ms0.copy(); ms1["diffuse_tex"] = 1; ms1["normal_tex"] = 2
    no_table = oracle.render((cam, sp, vs, ix, ms1), 40, 24, 1, 4, 1, 1, textures=[])
    assert_bits_equal(no_table, base, "handles but no table")
    textured = oracle.render((cam, sp, vs, ix, ms1), 40, 24, 1, 4, 1, 1, textures=scene.textures)
    assert not np.array_equal(textured, base)


def test_flat_normal_map_and_white_diffuse_are_identity(oracle):
    """n = (0, 0, 1) in tangent space keeps the interpolated normal's direction; a white texel
    keeps the colour.  (128 -> 2*128/255-1 != 0, so use an exactly representable map: 255 -> +1
    on z with x = y = 127.5 impossible in 8 bits; check instead that the perturbed image stays
    close while a white diffuse map is exact.)"""
    scene, sc = _cfg("c2_cube", 40, 24)
    cam, sp, vs, ix, ms = sc
    ms = ms.copy(); ms["normal_tex"] = 0
    white = np.full((4, 4, 4), 255, np.uint8)
    a = oracle.render((cam, sp, vs, ix, ms), 40, 24, 1, 4, 1, 1, textures=[white])
    ms2 = ms.copy(); ms2["diffuse_tex"] = 0
    b = oracle.render((cam, sp, vs, ix, ms2), 40, 24, 1, 4, 1, 1, textures=[white])
    assert_bits_equal(a, b, "white diffuse texture")      # byte 255 * f32(1/255) == 1.0f exactly


@pytest.mark.gpu
@pytest.mark.parametrize("w,h,frames,bounces", [(160, 90, 2, 8), (64, 64, 3, 2)])
def test_textured_cube_matches_oracle(rtlib, oracle, w, h, frames, bounces):
    scene, sc = _cfg("c2_cube", w, h)
    tr = rtlib.RayTracer(0, w, h)
    tr.upload_textures(scene.textures)              # order independent of the scene upload
    tr.upload_scene(*scene.arrays())
    tr.set_camera(scene.camera_uniform())
    tr.set_params(number_of_bounces=bounces, rays_per_pixel=1, skybox=1, frames=0)
    tr.set_option(rtlib._lib.RT_OPT_PRIMARY_IDS, 1)
    tr.render_frames(0, frames)
    img, ids, ctr = tr.read_accum(), tr.read_primary_ids(), tr.counters()
    ref_ids = np.full((h, w), 0xFFFFFFFF, np.uint32)
    ref_ctr = np.zeros(3, np.uint64)
    ref = oracle.render(sc, w, h, frames, bounces, 1, 1, 0, ref_ids, ref_ctr, textures=scene.textures)
    assert np.array_equal(ids, ref_ids) and ctr["rays"] == int(ref_ctr[0])
    assert_bits_equal(img, ref, "textured cube")
    # brute-force traversal gives the same textured image
    tr.set_option(rtlib._lib.RT_OPT_TRAVERSAL, 1); tr.clear(); tr.render_frames(0, frames)
    assert_bits_equal(tr.read_accum(), ref, "textured cube, exhaustive traversal")
    # dropping the table turns the handles into "none": the untextured image of the same records
    tr.set_option(rtlib._lib.RT_OPT_TRAVERSAL, 0); tr.upload_textures([]); tr.clear(); tr.render_frames(0, frames)
    assert_bits_equal(tr.read_accum(), oracle.render(sc, w, h, frames, bounces, 1, 1, 0), "handles without a table")
    tr.close()


@pytest.mark.gpu
def test_texture_upload_errors(rtlib):
    import ctypes as C
    tr = rtlib.RayTracer(0, 16, 16)
    with pytest.raises(ValueError):
        tr.upload_textures([np.zeros((4, 4, 3), np.uint8)])
    bad = (rtlib._lib.Texture * 1)(rtlib._lib.Texture(0, 4, None))
    rc = rtlib._lib.lib().rt_upload_textures(tr._h, C.cast(bad, C.c_void_p), 1)
    assert rc == rtlib._lib.RT_ERR_INVALID_ARGUMENT
    tr.upload_textures([np.zeros((3, 5, 4), np.uint8), np.zeros((1, 1, 4), np.uint8)])
    tr.upload_textures([])
    tr.close()
