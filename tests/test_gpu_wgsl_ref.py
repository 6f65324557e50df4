"""The CUDA path through the C ABI against oracle/_ref — the reference's own shader, translated
mechanically (oracle/make_ref.py) and compiled with the arithmetic contract — bit for bit; plus the
two full-size checks VERDICT r1 asked for (C1 at its BASELINE size, C5 records through a CPU oracle).

On the GPU box /root/reference does not exist: the prebuilt oracle/_ref libraries travel with the
snapshot.  Without them the `_ref` tests skip (the committed goldens, which are `_ref` outputs,
still run in test_goldens.py).
"""
import numpy as np
import pytest

from oracle import make_ref
from util import assert_bits_equal

pytestmark = pytest.mark.gpu
needs_ref = pytest.mark.skipif(make_ref.build() is None, reason="oracle/_ref not prebuilt")
MISS = 0xFFFFFFFF


def _gpu(rt, scene, w, h, frames, bounces, rpp=1, sky=0, first=0, seed_mode=0):
    cam, sp, vs, ix, ms = scene
    tr = rt.RayTracer(0, w, h)
    tr.upload_scene(sp, vs, ix, ms)
    tr.set_camera(cam.view(rt.CAMERA_DT) if cam.dtype != rt.CAMERA_DT else cam)
    tr.set_params(number_of_bounces=bounces, rays_per_pixel=rpp, skybox=sky, frames=first)
    tr.set_option(rt._lib.RT_OPT_PRIMARY_IDS, 1)
    if seed_mode:
        tr.set_option(rt._lib.RT_OPT_SEED_MODE, seed_mode)
    tr.render_frames(first, frames)
    img, ids, ctr = tr.read_accum(), tr.read_primary_ids(), tr.counters()
    tr.close()
    return img, ids, ctr


def _ref(scene, w, h, frames, bounces, rpp=1, sky=0, first=0, x_step=1, y_step=1):
    from oracle import wgsl_ref as wr
    ids = np.full((h, w), MISS, np.uint32)
    ctr = np.zeros(3, np.uint64)
    aov = wr.Aov(w, h)
    img = wr.render(scene, w, h, frames, bounces, rpp, sky, first, ids, ctr, x_step, y_step, contract=True, aov=aov)
    assert int(aov.mismatches[0]) == 0
    return img, ids, ctr, aov


def _builtin(name, w, h):
    from oracle import host_ref
    return getattr(host_ref, "scene_" + name)(np.float32(w) / np.float32(h))


def _config(name, w, h):
    from oracle import rt_oracle as ro
    from ray_tracer_b200 import configs
    t = configs.build(name, w, h)[0].as_tuple()
    return (t[0].view(ro.CAMERA_DT), t[1].view(ro.SPHERE_DT), t[2].view(ro.VERTEX_DT), t[3], t[4].view(ro.MESH_DT))


CASES = [
    # scene, w, h, frames, bounces, rpp, sky, first_frame
    ("balls", 96, 96, 3, 4, 1, 0, 0),
    ("balls", 60, 44, 2, 3, 1, 1, 1),            # start-up dispatch, not a multiple of 8
    ("metal", 96, 64, 2, 6, 2, 1, 0),            # glass, rpp = 2
    ("room", 96, 64, 2, 5, 1, 0, 0),             # triangles + spheres
    ("random_balls", 64, 40, 2, 4, 1, 1, 3),     # 487 spheres, thin lens
    ("balls", 40, 32, 1, 3, 0, 1, 0),            # rpp = 0
    ("room", 40, 32, 2, -1, 1, 1, 0),            # bounces = -1
    ("c2_cube_plain", 64, 36, 1, 8, 1, 1, 0),
    ("c3_teapot", 64, 36, 1, 8, 1, 1, 0),
    ("c4_tree", 64, 36, 1, 8, 1, 1, 0),
]


@needs_ref
@pytest.mark.parametrize("name,w,h,frames,bounces,rpp,sky,first", CASES)
def test_cuda_equals_translated_reference(rtlib, name, w, h, frames, bounces, rpp, sky, first):
    scene = _config(name, w, h) if name.startswith("c") else _builtin(name, w, h)
    gi, gid, gc = _gpu(rtlib, scene, w, h, frames, bounces, rpp, sky, first)
    ri, rid, rc, aov = _ref(scene, w, h, frames, bounces, rpp, sky, first)
    assert np.array_equal(gid, rid), "primary ids differ at %d pixels" % int((gid != rid).sum())
    assert gc["rays"] == int(rc[0]) == aov.calls("calculate_ray_collions")
    assert gc["samples"] == aov.calls("trace")
    assert_bits_equal(gi, ri, name, any_nan=(rpp == 0))


@needs_ref
def test_c1_at_baseline_size_equals_translated_reference(rtlib, oracle):
    """C1 as BASELINE.json states it: `balls`, 800x800, 16 spp = 8 dispatches, 4 bounces, sky off — every pixel, against
    the translated reference AND the hand oracle."""
    w = h = 800
    scene = _builtin("balls", w, h)
    gi, gid, gc = _gpu(rtlib, scene, w, h, 8, 4, 1, 0, 0)
    ri, rid, rc, aov = _ref(scene, w, h, 8, 4, 1, 0, 0)
    assert np.array_equal(gid, rid)
    assert gc["rays"] == int(rc[0]) and gc["samples"] == w * h * 8 * 2 == aov.calls("trace")
    assert_bits_equal(gi, ri, "c1 800x800 vs oracle/_ref")
    ids = np.full((h, w), MISS, np.uint32); ctr = np.zeros(3, np.uint64)
    ci = oracle.render(scene, w, h, 8, 4, 1, 0, 0, ids, ctr)
    assert np.array_equal(gid, ids) and gc["rays"] == int(ctr[0])
    assert_bits_equal(gi, ci, "c1 800x800 vs hand oracle")


def test_c5_records_through_the_cpu_oracle(rtlib, oracle):
    """C5 (100,000 spheres + 64 teapot instances aliasing one index range = 1,105,056 primitives): the CPU oracle's
    exhaustive scan on a strided pixel set (every 24th pixel of 480x270 in x and y: 240 pixels, 1 dispatch) against the
    GPU's hierarchy, bit for bit."""
    w, h, st = 480, 270, 24
    scene = _config("c5_stress", w, h)
    gi, gid, gc = _gpu(rtlib, scene, w, h, 1, 8, 1, 1, 0)
    ids = np.full((h, w), MISS, np.uint32); ctr = np.zeros(3, np.uint64)
    ci = oracle.render(scene, w, h, 1, 8, 1, 1, 0, ids, ctr, st, st)
    assert np.array_equal(gid[::st, ::st], ids[::st, ::st])
    assert (ids[::st, ::st] != MISS).sum() > 40          # the sample does see geometry
    assert_bits_equal(gi[::st, ::st], ci[::st, ::st], "c5 strided")


# ---- launch structure and seed modes ----------------------------------------------------------------
@pytest.mark.parametrize("name,w,h,frames,first", [("room", 96, 64, 5, 0), ("balls", 100, 76, 3, 1), ("c3_teapot", 160, 90, 4, 0)])
def test_item_modes_agree(rtlib, name, w, h, frames, first):
    """RT_OPT_ITEM_MODE 2 (one lane owns a pixel for all fused dispatches and applies wgsl:59-66 itself) == 1 ((pixel, dispatch)
    items, staged frags, blend kernel) == one launch per dispatch (auto: pixel items), bit for bit."""
    scene = _config(name, w, h) if name.startswith("c") else _builtin(name, w, h)
    cam, sp, vs, ix, ms = scene
    imgs = []
    for mode, fpl in ((2, 0), (1, 0), (0, 1), (1, 2), (2, 2)):
        tr = rtlib.RayTracer(0, w, h)
        tr.upload_scene(sp, vs, ix, ms)
        tr.set_camera(cam.view(rtlib.CAMERA_DT))
        tr.set_params(number_of_bounces=4, rays_per_pixel=1, skybox=1, frames=first)
        tr.set_option(rtlib._lib.RT_OPT_ITEM_MODE, mode)
        tr.set_option(rtlib._lib.RT_OPT_FRAMES_PER_LAUNCH, fpl)
        tr.render_frames(first, frames)
        imgs.append((tr.read_accum(), tr.counters()))
        tr.close()
    assert imgs[0][1]["launches"] == 1 and imgs[2][1]["launches"] == frames
    for img, ctr in imgs[1:]:
        assert ctr["rays"] == imgs[0][1]["rays"]
        assert_bits_equal(img, imgs[0][0], name)


def test_wide_seed_mode_at_4k(rtlib, oracle):
    """RT_OPT_SEED_MODE = 1: the seed's f32 -> u32 conversion wraps modulo 2^32 where strict mode saturates (3840x2160, rows
    >= 518).  CUDA == hand oracle (== oracle/_ref when present) on a strided pixel set in wide mode; wide == strict on the rows
    below 518 and differs above."""
    w, h, sx, sy = 3840, 2160, 96, 54
    scene = _builtin("balls", w, h)
    gw, _, _ = _gpu(rtlib, scene, w, h, 2, 3, 1, 1, 0, seed_mode=1)
    gs, _, _ = _gpu(rtlib, scene, w, h, 2, 3, 1, 1, 0, seed_mode=0)
    with oracle.seed_mode(1):
        cw = oracle.render(scene, w, h, 2, 3, 1, 1, 0, None, None, sx, sy)
        assert oracle.pixel_seed(1, 518, w, h, 0) == 0x177000 and oracle.pixel_seed(0, 517, w, h, 0) == 0xff98d000
    assert oracle.pixel_seed(1, 518, w, h, 0) == 0xFFFFFFFF
    assert_bits_equal(gw[::sy, ::sx], cw[::sy, ::sx], "wide seed mode, 4K strided")
    assert np.array_equal(gw[:516].view(np.uint32), gs[:516].view(np.uint32))     # rows 0..515: every seed is below 2^32 (row 517 crosses it at x = 1749)
    assert not np.array_equal(gw[520:].view(np.uint32), gs[520:].view(np.uint32))
    if make_ref.build() is not None:
        from oracle import wgsl_ref as wr
        with wr.seed_mode(1):
            rw = wr.render(scene, w, h, 2, 3, 1, 1, 0, None, None, sx, sy)
        assert_bits_equal(gw[::sy, ::sx], rw[::sy, ::sx], "wide seed mode vs oracle/_ref")


def test_pipelined_launches_are_result_invisible(rtlib):
    """A render call that needs several launches (dispatches beyond the staging budget) alternates them between two streams
    (RT_OPT_PIPELINE): same image and counters as serial launches and as one long launch, also when resumed and sharded."""
    w, h, frames = 240, 136, 41
    scene = _config("c3_teapot", w, h)
    cam, sp, vs, ix, ms = scene
    out = []
    for mib, pipe, shard in ((4096, 1, None), (16, 1, None), (16, 0, None), (16, 1, (1, 2, 1, 3)), (4096, 1, (1, 2, 1, 3))):
        tr = rtlib.RayTracer(0, w, h)
        tr.set_option(rtlib._lib.RT_OPT_STAGING_MIB, mib); tr.set_option(rtlib._lib.RT_OPT_PIPELINE, pipe)
        tr.upload_scene(sp, vs, ix, ms); tr.set_camera(cam.view(rtlib.CAMERA_DT))
        tr.set_params(number_of_bounces=4, rays_per_pixel=1, skybox=1, frames=0)
        if shard:
            tr.set_shard(*shard)
        tr.render_frames(0, frames)
        tr.render_frames(frames, 5)                        # a second call resumes the running mean
        out.append((tr.read_accum(), tr.counters()))
        tr.close()
    assert out[0][1]["launches"] == 2 and out[1][1]["launches"] >= 4 and out[2][1]["launches"] >= 3
    for k in (1, 2):
        assert out[k][1]["rays"] == out[0][1]["rays"]
        assert_bits_equal(out[k][0], out[0][0], "pipelined / serial launches")
    assert out[3][1]["rays"] == out[4][1]["rays"]
    assert_bits_equal(out[3][0], out[4][0], "pipelined launches under tile + sample sharding")


def test_learned_tile_order_is_result_invisible(rtlib):
    """RT_OPT_TILE_ORDER: from the second render call on the tiles that were heaviest in the previous call enter the queue first.
    Same image, counters and primary ids as raster order: over several calls, launch lengths, tile + sample shards, pipelined
    launches, single dispatches and a resize."""
    L = rtlib._lib
    out = {}
    for order in (0, 1):
        w, h = 200, 136
        scene = _config("c3_teapot", w, h)
        cam, sp, vs, ix, ms = scene
        tr = rtlib.RayTracer(0, w, h)
        tr.set_option(L.RT_OPT_TILE_ORDER, order); tr.set_option(L.RT_OPT_PRIMARY_IDS, 1)
        tr.upload_scene(sp, vs, ix, ms); tr.set_camera(cam.view(rtlib.CAMERA_DT))
        tr.set_params(number_of_bounces=4, rays_per_pixel=1, skybox=1, frames=0)
        res = []
        f = 0
        for n in (1, 3, 1, 20, 2):                                   # calls of different lengths resume the running mean
            tr.render_frames(f, n); f += n
            res.append(tr.read_accum())
        res.append(tr.read_primary_ids())
        tr.set_params(number_of_bounces=4, rays_per_pixel=1, skybox=1, frames=f)
        tr.dispatch(); res.append(tr.read_accum())                   # the reference's own single dispatch
        tr.set_shard(1, 3, 1, 2)                                     # another tile list: the order starts over
        tr.render_frames(0, 6); tr.render_frames(6, 6); res.append(tr.read_accum())
        tr.set_shard(0, 1, 0, 1)
        tr.set_option(L.RT_OPT_STAGING_MIB, 16); tr.render_frames(0, 60); tr.render_frames(60, 50); tr.render_frames(0, 40); tr.render_frames(40, 40); res.append(tr.read_accum())    # several (pipelined) launches per call
        res.append(tr.counters()["rays"])
        tr.update_bind_group(96, 72)                                 # resize
        scene2 = _config("c3_teapot", 96, 72)
        tr.set_camera(scene2[0].view(rtlib.CAMERA_DT)); tr.set_params(number_of_bounces=4, rays_per_pixel=1, skybox=1, frames=0, width=96, height=72)
        tr.render_frames(0, 2); tr.render_frames(2, 2); res.append(tr.read_accum())
        tr.close()
        out[order] = res
    for k, (a, b) in enumerate(zip(out[0], out[1])):
        if isinstance(a, np.ndarray) and a.dtype == np.float32:
            assert_bits_equal(b, a, "tile order, step %d" % k)
        else:
            assert np.array_equal(a, b), k
