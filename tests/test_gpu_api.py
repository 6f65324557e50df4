"""GPU tests of the rest of the C ABI: sharding options, Context frame logic (the C++ host
mirror), error behaviour, resize / write-back / tone-map read-out."""
import numpy as np
import pytest

from util import assert_bits_equal

pytestmark = pytest.mark.gpu
W, H, BOUNCES = 72, 40, 4


def _scene(name="room"):
    from oracle import host_ref as hr
    return getattr(hr, "scene_" + name)(np.float32(W) / np.float32(H))


def _tracer(rt, scene, **params):
    cam, sp, vs, ix, ms = scene
    tr = rt.RayTracer(0, W, H)
    tr.upload_scene(sp, vs, ix, ms)
    tr.set_camera(cam)
    p = dict(number_of_bounces=BOUNCES, rays_per_pixel=1, skybox=1, frames=0)
    p.update(params)
    tr.set_params(**p)
    return tr


def test_tile_sharding_is_a_bit_exact_partition(rtlib):
    from ray_tracer_b200 import sharding
    scene = _scene()
    tr = _tracer(rtlib, scene); tr.render_frames(0, 3); full = tr.read_accum(); tr.close()
    for world in (2, 3):
        total = np.zeros_like(full)
        owner = sharding.tile_owner(W, H, world)
        for r in range(world):
            tr = _tracer(rtlib, scene)
            tr.set_shard(tile_rank=r, tile_world=world)
            tr.render_frames(0, 3)
            part = tr.read_accum(); tr.close()
            assert not part[owner != r].any(), "rank %d wrote outside its tiles" % r
            assert_bits_equal(part[owner == r], full[owner == r], "tiles of rank %d/%d" % (r, world))
            total += part
        assert np.array_equal(total, full)


def test_sample_sharding_matches_per_rank_running_means(rtlib):
    from ray_tracer_b200 import sharding
    scene = _scene("balls")
    frames, world = 6, 2
    tr = _tracer(rtlib, scene); tr.render_frames(0, frames); full = tr.read_accum(); tr.close()
    # frag of dispatch f alone: frames=-f overwrites and seeds with |frames| (wgsl:64-66,308)
    frags = []
    for f in range(frames):
        tr = _tracer(rtlib, scene, frames=-f); tr.dispatch(); frags.append(tr.read_accum()); tr.close()
    combined = np.zeros_like(full, dtype=np.float64)
    for r in range(world):
        tr = _tracer(rtlib, scene)
        tr.set_shard(frame_rank=r, frame_world=world)
        tr.render_frames(0, frames)
        got = tr.read_accum(); n = tr.counters()["dispatches"]; tr.close()
        own = sharding.frames_of_rank(0, frames, r, world)
        assert n == len(own)
        acc = None
        for f in own:
            k = f // world
            if k >= 1:
                w = np.float32(1.0) / np.float32(k + 1)
                acc = (acc * (np.float32(1.0) - w) + frags[f] * w).astype(np.float32)
            else:
                acc = frags[f]
        assert_bits_equal(got, acc, "rank %d running mean" % r)
        combined += got.astype(np.float64) * len(own) / frames
    assert np.allclose(combined, full, rtol=2e-6, atol=1e-6)


def test_dispatch_equals_render_frames_and_continuation(rtlib):
    scene = _scene()
    a = _tracer(rtlib, scene); a.render_frames(0, 5); ia = a.read_accum(); a.close()
    b = _tracer(rtlib, scene)
    for f in range(5):
        b.set_params(number_of_bounces=BOUNCES, rays_per_pixel=1, skybox=1, frames=f)
        b.dispatch()
    ib = b.read_accum(); b.close()
    assert_bits_equal(ia, ib, "5 dispatches vs render_frames(0,5)")
    c = _tracer(rtlib, scene); c.render_frames(0, 2); c.render_frames(2, 3); ic = c.read_accum(); c.close()
    assert_bits_equal(ia, ic, "render_frames continued across calls")


def test_context_frame_logic_matches_reference_semantics(rtlib, oracle):
    """Context::new/update/render/clear_accululation/resize (context.rs:86-94,126-146,176-208)."""
    from ray_tracer_b200 import host
    from oracle import host_ref as hr
    ctx = host.Context(0, W, H)
    p = ctx.params
    assert (p["number_of_bounces"][0], p["rays_per_pixel"][0], p["skybox"][0], p["frames"][0], p["accumulate"][0]) == (3, 1, 0, 0, 1)
    scene = hr.scene_balls(np.float32(W) / np.float32(H))             # Context::new loads Scene::balls
    # first redraw after start-up: update bumps frames to 1 -> blend 50/50 with the zero texture
    ctx.update(); ctx.render()
    assert ctx.params["frames"][0] == 1
    want = oracle.render(scene, W, H, 1, 3, 1, 0, 1)
    assert_bits_equal(ctx.read_texture(), want, "start-up dispatch")
    ctx.run_frames(3)                                                  # frames 2,3,4, fused
    want = oracle.render(scene, W, H, 3, 3, 1, 0, 2, accum=want)
    assert_bits_equal(ctx.read_texture(), want, "run_frames")
    ctx.clear_accululation()                                           # frames=-1 -> next update 0 -> overwrite
    ctx.update(); ctx.render(); ctx.update(); ctx.render()
    assert ctx.params["frames"][0] == 1
    assert_bits_equal(ctx.read_texture(), oracle.render(scene, W, H, 2, 3, 1, 0, 0), "after clear")
    ctx.set(accumulate=0)                                              # frames pinned to -1: static noise
    ctx.update(); ctx.render(); one = ctx.read_texture(); ctx.update(); ctx.render()
    assert ctx.params["frames"][0] == -1
    assert_bits_equal(ctx.read_texture(), one, "accumulate off")
    assert_bits_equal(one, oracle.render(scene, W, H, 1, 3, 1, 0, -1), "accumulate off image")
    ctx.set(accumulate=1)
    ctx.resize(48, 32)                                                 # new zero texture, frames=-1, aspect updated
    assert not ctx.read_texture().any() and ctx.params["frames"][0] == -1
    ctx.update(); ctx.render()
    assert_bits_equal(ctx.read_texture(), oracle.render(hr.scene_balls(np.float32(48) / np.float32(32)), 48, 32, 1, 3, 1, 0, 0), "after resize")
    ctx.set_scene(2)                                                   # room (context.rs:270-272)
    ctx.update(); ctx.render()
    assert_bits_equal(ctx.read_texture(), oracle.render(hr.scene_room(np.float32(48) / np.float32(32)), 48, 32, 1, 3, 1, 0, 0), "scene switch")
    ctx.close()


def test_error_behaviour(rtlib):
    L, E = rtlib.RayTracer, rtlib.RtError
    tr = L(0, 16, 16)
    with pytest.raises(E) as e:
        tr.dispatch()
    assert e.value.code == rtlib._lib.RT_ERR_NO_SCENE
    cam, sp, vs, ix, ms = _scene()
    bad = ms.copy(); bad["first"][0] = 40                              # reads past the index array
    with pytest.raises(E) as e:
        tr.upload_scene(sp, vs, ix, bad)
    assert e.value.code == rtlib._lib.RT_ERR_BAD_INDEX
    bad_ix = ix.copy(); bad_ix[3] = 999
    with pytest.raises(E) as e:
        tr.upload_scene(sp, vs, bad_ix, ms)
    assert e.value.code == rtlib._lib.RT_ERR_BAD_INDEX and "vertex" in str(e.value)
    tr.upload_scene(sp, vs, ix, ms)
    with pytest.raises(E):
        tr.set_params(width=32, height=16, number_of_bounces=1, rays_per_pixel=1)   # size mismatch
    with pytest.raises(E):
        tr.set_shard(tile_rank=2, tile_world=2)
    with pytest.raises(E):
        tr.set_option(rtlib._lib.RT_OPT_LEAF_SIZE, 9)
    with pytest.raises(E):
        L(99, 16, 16)                                                  # no such device
    tr.close()


def test_resize_write_back_and_tonemap(rtlib):
    scene = _scene("balls")
    tr = _tracer(rtlib, scene)
    tr.render_frames(0, 2)
    img = tr.read_accum()
    # checkpoint / resume: saved image + continuing frames == uninterrupted run
    tr2 = _tracer(rtlib, scene); tr2.write_accum(img); tr2.render_frames(2, 2); resumed = tr2.read_accum(); tr2.close()
    tr.render_frames(2, 2)
    assert_bits_equal(tr.read_accum(), resumed, "resume from a saved accumulation image")
    # what the sRGB swapchain showed: clamp + OETF, 8 bits (render.wgsl:34-39, context.rs:75-80)
    rgba = tr.read_rgba8()
    lin = np.clip(np.nan_to_num(tr.read_accum()[..., :3], nan=0.0), 0, 1).astype(np.float64)
    srgb = np.where(lin <= 0.0031308, 12.92 * lin, 1.055 * lin ** (1 / 2.4) - 0.055)
    assert np.abs(rgba[..., :3].astype(np.int32) - np.rint(srgb * 255).astype(np.int32)).max() <= 1
    tr.update_bind_group(40, 24)                                       # resize: zero image
    assert tr.read_accum().shape == (24, 40, 4) and not tr.read_accum().any()
    tr.clear()
    tr.close()


def test_counters_and_bvh_stats(rtlib):
    scene = _scene()
    tr = _tracer(rtlib, scene)
    tr.set_option(rtlib._lib.RT_OPT_DETAIL_COUNTERS, 1)
    tr.render_frames(0, 2)
    c, st = tr.counters(), tr.bvh_stats()
    assert c["samples"] == W * H * 2 * 2 and c["rays"] >= c["samples"] and c["hit_segments"] <= c["rays"]
    assert c["node_visits"] > 0 and c["tri_tests"] > 0 and c["sphere_tests"] > 0 and c["dispatches"] == 2 and c["kernel_ms"] > 0
    assert st["n_prims"] == 2 + 14 and st["max_depth"] >= 2
    tr.reset_counters()
    assert tr.counters()["rays"] == 0
    tr.close()


def test_update_records_equals_fresh_upload(rtlib):
    """rt_update_records (refit on the kept topology) == rt_upload_scene of the moved records,
    bit for bit, for both builders; topology changes are rejected."""
    from ray_tracer_b200 import configs
    scene, meta = configs.build("c5_stress", 96, 54)
    cam, sp, vs, ix, ms = scene.as_tuple()
    sp, ms = sp[:2000].copy(), ms.copy()
    rng = np.random.default_rng(11)
    sp2, ms2 = sp.copy(), ms.copy()
    ms2["pos"] += rng.normal(0, 40.0, ms2["pos"].shape).astype(np.float32)       # the teapot instances drift
    sp2["position"] += rng.normal(0, 25.0, sp2["position"].shape).astype(np.float32)
    sp2["radius"] *= np.float32(1.5)
    ms2["color"][:, 0] = 0.9
    for builder, inst in ((0, 1), (1, 1), (0, 0)):               # instanced (a top-level refit only) and flattened records
        def render(tr):
            tr.set_camera(cam); tr.set_params(number_of_bounces=3, rays_per_pixel=1, skybox=1, frames=0)
            tr.set_option(rtlib._lib.RT_OPT_PRIMARY_IDS, 1)
            tr.clear(); tr.reset_counters(); tr.render_frames(0, 1)
            return tr.read_accum(), tr.read_primary_ids(), tr.counters()["rays"]
        a = rtlib.RayTracer(0, 96, 54); a.set_option(rtlib._lib.RT_OPT_BUILDER, builder); a.set_option(rtlib._lib.RT_OPT_INSTANCING, inst)
        a.upload_scene(sp, vs, ix, ms)
        assert a.bvh_stats()["n_prims"] == (2000 + 128 if inst else 2000 + 64 * 15704)
        img0, _, _ = render(a)
        a.update_records(sp2, ms2)
        img_a, ids_a, rays_a = render(a)
        b = rtlib.RayTracer(0, 96, 54); b.set_option(rtlib._lib.RT_OPT_BUILDER, builder); b.set_option(rtlib._lib.RT_OPT_INSTANCING, 0)
        b.upload_scene(sp2, vs, ix, ms2)                                           # the fresh upload is the flattened one
        img_b, ids_b, rays_b = render(b)
        assert not np.array_equal(img0, img_a)
        assert np.array_equal(ids_a, ids_b) and rays_a == rays_b
        assert np.array_equal(img_a.view(np.uint32), img_b.view(np.uint32))
        a.update_records(sp, ms)                                                   # and back
        assert np.array_equal(render(a)[0].view(np.uint32), img0.view(np.uint32))
        bad = ms2.copy(); bad["triangles"][0] -= 1
        with pytest.raises(rtlib.RtError):
            a.update_records(sp2, bad)
        with pytest.raises(rtlib.RtError):
            a.update_records(sp2[:-1], ms2)
        a.close(); b.close()


def test_wavefront_kernels_respect_the_shard(rtlib):
    """RT_OPT_KERNEL = 1 under tile + sample sharding, an odd image size and a resumed accumulation: same image as
    the megakernel with the same shard."""
    scene = _scene()
    out = []
    for kernel in (0, 1):
        cam, sp, vs, ix, ms = scene
        tr = rtlib.RayTracer(0, 75, 43)
        tr.set_option(rtlib._lib.RT_OPT_KERNEL, kernel)
        tr.set_option(rtlib._lib.RT_OPT_WAVE_SLOTS, 2048)
        tr.upload_scene(sp, vs, ix, ms); tr.set_camera(cam)
        tr.set_params(number_of_bounces=BOUNCES, rays_per_pixel=2, skybox=1, frames=0)
        tr.set_shard(tile_rank=1, tile_world=3, frame_rank=1, frame_world=2)
        tr.render_frames(1, 3)
        tr.render_frames(7, 2)
        out.append((tr.read_accum(), tr.counters()))
        tr.close()
    assert out[0][1]["rays"] == out[1][1]["rays"] and out[0][1]["samples"] == out[1][1]["samples"]
    assert_bits_equal(out[0][0], out[1][0], "wavefront under sharding")
