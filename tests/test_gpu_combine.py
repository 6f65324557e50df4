"""GPU tests of rt_combine_peers / rt_combine_peers_rgba8: the shards' images folded into one by a single kernel
(over NVLink peer memory when the contexts live on different devices; the same kernel on one device otherwise).
No reference counterpart (upstream is single-device): the checks are against the single-context image, bit for
bit for tile shards, and against the weighted sum evaluated in numpy for sample shards."""
import ctypes as C

import numpy as np
import pytest

from util import assert_bits_equal

pytestmark = pytest.mark.gpu
W, H, BOUNCES = 88, 56, 4


def _devices(n):
    """n device indices: distinct GPUs when the box has them, else all on device 0."""
    import torch
    have = torch.cuda.device_count()
    return [k % max(have, 1) for k in range(n)]


def _tracer(rt, dev, scene, **shard):
    cam, sp, vs, ix, ms = scene
    tr = rt.RayTracer(dev, W, H)
    tr.upload_scene(sp, vs, ix, ms)
    tr.set_camera(cam)
    tr.set_params(number_of_bounces=BOUNCES, rays_per_pixel=1, skybox=1, frames=0)
    tr.set_shard(**shard)
    return tr


def _scene(name):
    from oracle import host_ref as hr
    return getattr(hr, "scene_" + name)(np.float32(W) / np.float32(H))


def test_tile_shards_combine_to_the_single_context_image_bit_for_bit(rtlib):
    scene, frames = _scene("room"), 4
    one = _tracer(rtlib, 0, scene); one.render_frames(0, frames); full = one.read_accum(); full8 = one.read_rgba8(); one.close()
    for world in (2, 3):
        trs = [_tracer(rtlib, d, scene, tile_rank=r, tile_world=world) for r, d in enumerate(_devices(world))]
        for t in trs:
            t.render_frames(0, frames)                     # asynchronous on every context's stream
        got8 = rtlib.combine_peers_rgba8(trs)               # fused combine + tone-map, accumulation images untouched
        assert np.array_equal(got8, full8)
        part0 = trs[0].read_accum()
        assert not np.array_equal(part0, full)
        rtlib.combine_peers(trs)
        assert_bits_equal(trs[0].read_combined(), full, "tile shards, world %d" % world)
        assert_bits_equal(trs[0].read_accum(), part0, "a combine leaves the accumulation image alone")
        # every context keeps its own tiles and can go on rendering; a second combine follows the single context
        for t in trs:
            t.render_frames(frames, 2)
        rtlib.combine_peers(trs)
        one = _tracer(rtlib, 0, scene); one.render_frames(0, frames + 2); full2 = one.read_accum(); one.close()
        assert_bits_equal(trs[0].read_combined(), full2, "tile shards after more frames, world %d" % world)
        for t in trs:
            t.close()


def test_tile_and_sample_shards_combine_like_the_weighted_sum(rtlib):
    scene, frames = _scene("balls"), 6
    T, F = 2, 2
    trs = [_tracer(rtlib, d, scene, tile_rank=k % T, tile_world=T, frame_rank=k // T, frame_world=F)
           for k, d in enumerate(_devices(T * F))]
    for t in trs:
        t.render_frames(0, frames)
    parts = [t.read_accum() for t in trs]
    from ray_tracer_b200 import sharding
    owner = sharding.tile_owner(W, H, T)
    want = np.zeros_like(parts[0])
    seen = np.zeros((H, W), bool)
    half = np.float32(1.0) / np.float32(F)
    for k, p in enumerate(parts):                           # array order, separately rounded multiply and add
        m = owner == k % T
        term = (p * half).astype(np.float32)
        want[m] = np.where(seen[m][:, None], (want[m] + term[m]).astype(np.float32), term[m])
        seen |= m
    rtlib.combine_peers(trs)
    got = trs[0].read_combined()
    assert [t.mean_count for t in trs] == [3, 3, 3, 3]
    assert_bits_equal(got, want, "2 tile x 2 sample shards")
    one = _tracer(rtlib, 0, scene); one.render_frames(0, frames); full = one.read_accum(); one.close()
    assert np.allclose(got, full, rtol=2e-6, atol=1e-6)     # equal to the sequential running mean up to f32 rounding
    # explicit weights: 3 frames of shard 0 against 1 of shard 1
    for t in trs:
        t.close()
    a = _tracer(rtlib, 0, scene); a.render_frames(0, 3)
    b = _tracer(rtlib, _devices(2)[1], scene); b.set_params(number_of_bounces=BOUNCES, rays_per_pixel=1, skybox=1, frames=-3); b.dispatch()
    pa, pb = a.read_accum(), b.read_accum()
    rtlib.combine_peers([a, b], weights=[0.75, 0.25])
    assert_bits_equal(a.read_combined(), ((pa * np.float32(0.75)).astype(np.float32) + (pb * np.float32(0.25)).astype(np.float32)).astype(np.float32), "weighted")
    one = _tracer(rtlib, 0, scene); one.render_frames(0, 4); full4 = one.read_accum(); one.close()
    assert np.allclose(a.read_combined(), full4, rtol=2e-6, atol=1e-6)
    a.close(); b.close()


def test_sample_shards_keep_accumulating_across_combines_and_uneven_counts(rtlib):
    """a combine must not disturb the shards' running means (render, combine, render, combine), and shards
    that hold different numbers of dispatches (n_frames not divisible by the world) enter with count weights."""
    scene = _scene("room")
    devs = _devices(3)
    trs = [_tracer(rtlib, d, scene, frame_rank=k, frame_world=3) for k, d in enumerate(devs)]
    for t in trs:
        t.render_frames(0, 7)                              # ranks hold 3, 2, 2 dispatches
    assert [t.mean_count for t in trs] == [3, 2, 2]
    parts = [t.read_accum() for t in trs]
    rtlib.combine_peers(trs)
    got = trs[0].read_combined()
    want = None
    for p, n in zip(parts, (3, 2, 2)):
        term = (p * (np.float32(n) / np.float32(7))).astype(np.float32)
        want = term if want is None else (want + term).astype(np.float32)
    assert_bits_equal(got, want, "count-weighted 3/2/2")
    one = _tracer(rtlib, 0, scene); one.render_frames(0, 7); full7 = one.read_accum()
    assert np.allclose(got, full7, rtol=3e-6, atol=2e-6)
    for t in trs:
        t.render_frames(7, 5)                              # frames 7..11 on top: 4 dispatches per rank now
    assert [t.mean_count for t in trs] == [4, 4, 4]
    rtlib.combine_peers(trs)
    one.render_frames(7, 5); full12 = one.read_accum(); one.close()
    assert np.allclose(trs[0].read_combined(), full12, rtol=3e-6, atol=2e-6)
    for t in trs:
        t.close()


def test_combine_errors(rtlib):
    scene = _scene("balls")
    a = _tracer(rtlib, 0, scene)
    b = rtlib.RayTracer(0, W + 8, H)
    with pytest.raises(rtlib.RtError):
        rtlib.combine_peers([a, b])                         # sizes differ
    with pytest.raises(rtlib.RtError):
        rtlib.combine_peers([a, a])                         # listed twice
    with pytest.raises(rtlib.RtError):
        rtlib.combine_peers([a] + [b] * 16)                 # more than 16
    assert rtlib._lib.lib().rt_combine_peers(None, 0, None) == rtlib._lib.RT_ERR_INVALID_ARGUMENT
    rtlib.combine_peers([a])                                # one context: weight 1, image unchanged
    a.render_frames(0, 2); img = a.read_accum()
    rtlib.combine_peers([a])
    assert_bits_equal(a.read_combined(), img, "single context")
    assert_bits_equal(a.read_accum(), img, "single context, accum")
    a.close(); b.close()


def test_context_mirror_shards_and_combine(rtlib):
    """The C++ Context mirror driven as two tile shards (same scene, same update / run_frames calls on both) and
    Context.combine: the image equals the single Context's bit for bit (context.rs frame logic unchanged per shard)."""
    from ray_tracer_b200 import host
    devs = _devices(2)
    one = host.Context(0, W, H)
    one.set(number_of_bounces=BOUNCES, skybox=1)
    one.set_scene(2)                                        # room (context.rs:261-278)
    one.run_frames(5); one.sync()
    full = one.read_texture(); frames_one = int(one.params["frames"][0]); one.close()
    shards = []
    for r in range(2):
        c = host.Context(devs[r], W, H)
        c.set(number_of_bounces=BOUNCES, skybox=1)
        c.set_scene(2)
        c.set_shard(tile_rank=r, tile_world=2)
        shards.append(c)
    for c in shards:
        c.run_frames(5)
    host.Context.combine(shards)
    got = shards[0].read_texture()
    assert int(shards[0].params["frames"][0]) == frames_one
    assert_bits_equal(got, full, "two Context shards combined")
    for c in shards:
        c.close()


def test_headless_render_driver(rtlib, tmp_path):
    """python -m ray_tracer_b200.render: a built-in scene on one context and as two tile shards gives the same PNG;
    a model file with a two-position turntable writes one image per position."""
    from ray_tracer_b200 import host, render
    a, b = str(tmp_path / "one.png"), str(tmp_path / "two.png")
    render.main(["--scene", "room", "--size", str(W), str(H), "--frames", "3", "--sky", "--out", a])
    render.main(["--scene", "room", "--size", str(W), str(H), "--frames", "3", "--sky", "--gpus", "2" if len(set(_devices(2))) == 2 else "1", "--out", b])
    ia, ib = host.load_png(a), host.load_png(b)
    assert ia.shape == (H, W, 4) and np.array_equal(ia, ib) and ia[..., :3].max() > 0
    obj = tmp_path / "quad.obj"
    obj.write_text("v -1 -1 0\nv 1 -1 0\nv 1 1 0\nv -1 1 0\nvn 0 0 1\nf 1//1 2//1 3//1\nf 1//1 3//1 4//1\n")
    out = render.main(["--model", str(obj), "--size", "64", "48", "--frames", "2", "--turntable", "40", "--positions", "2",
                       "--out", str(tmp_path / "t.png")])
    imgs = [host.load_png(o["image"]) for o in out]
    assert len(imgs) == 2 and imgs[0].shape == (48, 64, 4) and not np.array_equal(imgs[0], imgs[1])
