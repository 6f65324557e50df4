"""GPU parity proper: the CUDA path through the C ABI vs the CPU oracle, bit for bit.

The arithmetic contract fixes every f32 operation that can change a pixel, so
the bar here is exact equality of the whole RGBA32F accumulation image, of the primary-hit
ids and of the ray (segment) count — not a tolerance.
"""
import numpy as np
import pytest

from util import assert_bits_equal

pytestmark = pytest.mark.gpu


def _render_gpu(rt, scene, w, h, frames, bounces, rpp=1, toggle=0, first_frame=0, brute=False, fpl=None):
    cam, sp, vs, ix, ms = scene
    tr = rt.RayTracer(0, w, h)
    tr.upload_scene(sp, vs, ix, ms)
    tr.set_camera(cam.view(rt.CAMERA_DT) if cam.dtype != rt.CAMERA_DT else cam)
    tr.set_params(number_of_bounces=bounces, rays_per_pixel=rpp, skybox=toggle, frames=first_frame)
    tr.set_option(rt._lib.RT_OPT_PRIMARY_IDS, 1)
    tr.set_option(rt._lib.RT_OPT_TRAVERSAL, 1 if brute else 0)
    if fpl is not None:
        tr.set_option(rt._lib.RT_OPT_FRAMES_PER_LAUNCH, fpl)
    tr.render_frames(first_frame, frames)
    img, ids, ctr = tr.read_accum(), tr.read_primary_ids(), tr.counters()
    tr.close()
    return img, ids, ctr


def _render_cpu(oracle, scene, w, h, frames, bounces, rpp=1, toggle=0, first_frame=0):
    ids = np.full((h, w), 0xFFFFFFFF, np.uint32)
    ctr = np.zeros(3, np.uint64)
    img = oracle.render(scene, w, h, frames, bounces, rpp, toggle, first_frame, ids, ctr)
    return img, ids, ctr


CASES = [
    # scene, w, h, frames, bounces, rpp, toggle, first_frame
    ("balls", 128, 128, 4, 4, 1, 0, 0),
    ("balls", 100, 76, 2, 3, 1, 1, 0),       # not a multiple of 8, sky on
    ("balls", 64, 64, 1, 3, 1, 0, 1),        # start-up dispatch: frames=1 blends with the zero image
    ("balls", 64, 64, 1, 3, 1, 0, -1),       # accumulate off / reset: frames=-1 overwrites
    ("balls", 64, 64, 2, 0, 3, 1, 0),        # bounces=0, rpp=3
    ("metal", 128, 96, 3, 6, 1, 1, 0),       # glass sphere
    ("room", 128, 128, 3, 4, 1, 0, 0),       # triangles + spheres, emissive quad
    ("room", 96, 64, 2, 8, 2, 1, 5),
    ("random_balls", 96, 64, 2, 4, 1, 1, 0), # 487 spheres, thin lens
    ("balls", 48, 40, 2, 3, 0, 1, 0),        # rays_per_pixel = 0: one sample divided by 0 (inf / NaN, wgsl:326)
    ("balls", 48, 40, 1, 3, -1, 1, 0),       # rays_per_pixel = -1: the sample loop never runs, 0 / -1
    ("room", 48, 40, 2, -1, 1, 1, 0),        # number_of_bounces = -1: paths end untraced, RNG still advances per sample
    ("balls", 3840, 2160, 1, 1, 1, 1, 0),    # 4K: the f32 seed saturates from row 518 on (strict mode)
]


def _scene(name, w, h):
    from oracle import host_ref
    return getattr(host_ref, "scene_" + name)(np.float32(w) / np.float32(h))


@pytest.mark.parametrize("name,w,h,frames,bounces,rpp,toggle,first", CASES)
def test_image_ids_rays_bit_exact(rtlib, oracle, name, w, h, frames, bounces, rpp, toggle, first):
    scene = _scene(name, w, h)
    gi, gid, gc = _render_gpu(rtlib, scene, w, h, frames, bounces, rpp, toggle, first)
    ci, cid, cc = _render_cpu(oracle, scene, w, h, frames, bounces, rpp, toggle, first)
    assert np.array_equal(gid, cid), "primary ids differ at %d pixels" % int((gid != cid).sum())
    assert gc["rays"] == int(cc[0])
    assert gc["samples"] == w * h * frames * (rpp + 1)
    assert_bits_equal(gi, ci, name, any_nan=(rpp == 0))     # rpp = 0 divides by zero: 0/0 is NaN on both sides, payload unspecified


@pytest.mark.parametrize("name", ["balls", "room", "random_balls"])
def test_bvh_equals_brute_force(rtlib, name):
    w, h = 96, 64
    scene = _scene(name, w, h)
    a, aid, ac = _render_gpu(rtlib, scene, w, h, 3, 5, 1, 1, 0, brute=False)
    b, bid, bc = _render_gpu(rtlib, scene, w, h, 3, 5, 1, 1, 0, brute=True)
    assert np.array_equal(aid, bid)
    assert ac["rays"] == bc["rays"]
    assert_bits_equal(a, b, name)


def test_fused_launch_equals_one_launch_per_dispatch(rtlib):
    w, h = 64, 64
    scene = _scene("room", w, h)
    a, _, ca = _render_gpu(rtlib, scene, w, h, 6, 4, fpl=64)
    b, _, cb = _render_gpu(rtlib, scene, w, h, 6, 4, fpl=1)
    assert ca["launches"] == 1 and cb["launches"] == 6
    assert_bits_equal(a, b, "fused vs per-dispatch launches")


def test_empty_scene_and_sky(rtlib, oracle):
    from oracle import host_ref
    w, h = 64, 48
    cam = host_ref.scene_balls(np.float32(w) / np.float32(h))[0]
    empty = (cam, np.zeros(0, oracle.SPHERE_DT), np.zeros(0, oracle.VERTEX_DT), np.zeros(0, np.uint32),
             np.zeros(0, oracle.MESH_DT))
    gi, gid, gc = _render_gpu(rtlib, empty, w, h, 2, 3, 1, 1, 0)
    ci, cid, cc = _render_cpu(oracle, empty, w, h, 2, 3, 1, 1, 0)
    assert (gid == 0xFFFFFFFF).all() and gc["rays"] == int(cc[0])
    assert_bits_equal(gi, ci, "empty scene")
