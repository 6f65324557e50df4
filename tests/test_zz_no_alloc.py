"""rt_render_frames_static — the render call that allocates nothing (scratch: the library's static device array of
RT_STATIC_SCRATCH_BYTES, rt_trace.cu) — against rt_render_frames, through the C ABI, bit for bit (the tolerance of every
integer / f32-contract test here), on two calls in a row and across contexts that share the array.

No reference counterpart.  This file sorts last on purpose.
It passes alone in a fresh process (the first call resolves the array's address) and after the others.
"""
import numpy as np
import pytest

from util import assert_bits_equal

pytestmark = pytest.mark.gpu
SCRATCH = 32 << 20


def _scene(name, w, h):
    if name.startswith("c"):
        from oracle import rt_oracle as ro
        from ray_tracer_b200 import configs
        t = configs.build(name, w, h)[0].as_tuple()
        return (t[0].view(ro.CAMERA_DT), t[1].view(ro.SPHERE_DT), t[2].view(ro.VERTEX_DT), t[3], t[4].view(ro.MESH_DT))
    from oracle import host_ref
    return getattr(host_ref, "scene_" + name)(np.float32(w) / np.float32(h))


def _ctx(rt, scene, w, h, bounces=4, rpp=1, sky=1):
    cam, sp, vs, ix, ms = scene
    tr = rt.RayTracer(0, w, h)
    tr.upload_scene(sp, vs, ix, ms)
    tr.set_camera(cam.view(rt.CAMERA_DT))
    tr.set_params(number_of_bounces=bounces, rays_per_pixel=rpp, skybox=sky, frames=0)
    tr.set_option(rt._lib.RT_OPT_PRIMARY_IDS, 1)
    return tr


def _static(tr, first, n):
    """The entry point itself, through ctypes; returns (status, message)."""
    rc = tr._L.rt_render_frames_static(tr._h, int(first), int(n))
    return rc, tr._L.rt_last_error(tr._h).decode()


@pytest.mark.parametrize("name,w,h", [("room", 160, 90), ("c3_teapot", 200, 120), ("random_balls", 100, 76)])
def test_static_scratch_render_equals_render_frames_twice(rtlib, name, w, h):
    scene = _scene(name, w, h)
    a, b = _ctx(rtlib, scene, w, h), _ctx(rtlib, scene, w, h)
    try:
        # call 1: dispatches 0..4 in one launch; call 2 resumes the running mean with 5..7 on the same (dirty) scratch
        for k, (first, n) in enumerate(((0, 5), (5, 3))):
            rc, msg = _static(a, first, n)
            assert rc == 0, msg
            b.render_frames(first, n)
            ia, ib = a.read_accum(), b.read_accum()
            assert_bits_equal(ia, ib, "%s, call %d" % (name, k + 1))
            ca, cb = a.counters(), b.counters()
            assert ca["rays"] == cb["rays"] and ca["samples"] == cb["samples"] and ca["dispatches"] == cb["dispatches"]
            if k == 0:
                assert np.array_equal(a.read_primary_ids(), b.read_primary_ids())
        assert a.counters()["launches"] == 2 and a.mean_count == b.mean_count == 8
        assert ia[..., :3].any()                                              # an image, not two untouched buffers (the sky may hold inf, as upstream)
    finally:
        a.close(); b.close()


def test_static_scratch_is_shared_between_contexts_and_sizes(rtlib):
    """Two contexts of different image sizes alternate on the one array (serialised by rt_sync): what a call finds in it
    — the other context's planes and a used queue head — must not show."""
    s1, s2 = _scene("balls", 96, 96), _scene("room", 64, 40)
    a1, b1 = _ctx(rtlib, s1, 96, 96, sky=0), _ctx(rtlib, s1, 96, 96, sky=0)
    a2, b2 = _ctx(rtlib, s2, 64, 40), _ctx(rtlib, s2, 64, 40)
    try:
        a1.set_shard(1, 3, 0, 1); b1.set_shard(1, 3, 0, 1)                   # tile sharding is supported
        for first, n in ((0, 2), (2, 4), (6, 1)):
            for a, b in ((a1, b1), (a2, b2)):
                rc, msg = _static(a, first, n)
                assert rc == 0, msg
                a.sync()
                b.render_frames(first, n)
                assert_bits_equal(a.read_accum(), b.read_accum(), "frames %d..%d" % (first, first + n - 1))
    finally:
        for t in (a1, b1, a2, b2):
            t.close()


def test_static_scratch_rejects_what_does_not_fit(rtlib):
    L = rtlib._lib
    w, h = 1920, 1080
    scene = _scene("balls", w, h)
    tr = _ctx(rtlib, scene, w, h, sky=0)
    try:
        assert 1 * w * h * 16 <= SCRATCH < 2 * w * h * 16
        rc, msg = _static(tr, 0, 2)                                          # two 1080p planes: 66 MB > 32 MiB
        assert rc == L.RT_ERR_INVALID_ARGUMENT and "static scratch" in msg and "at most 1 dispatches" in msg
        assert tr.counters()["launches"] == 0 and not tr.read_accum().any()  # nothing was launched
        rc, msg = _static(tr, 0, 1)                                          # one plane fits
        assert rc == 0, msg
        ref = _ctx(rtlib, scene, w, h, sky=0)
        ref.render_frames(0, 1)
        assert_bits_equal(tr.read_accum(), ref.read_accum(), "1080p, one dispatch")
        ref.close()
        tr.set_shard(0, 1, 1, 2)                                             # sample sharding: refused with a message
        rc, msg = _static(tr, 0, 1)
        assert rc == L.RT_ERR_UNSUPPORTED and "frame_world" in msg
    finally:
        tr.close()
