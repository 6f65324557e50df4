"""rt_comm_init / rt_combine / rt_combine_rgba8: the multi-PROCESS combine (one process per GPU, NCCL behind the C ABI).
Needs two or more GPUs on the box (skipped otherwise).  The ranks are spawned here, get the communicator id from the
parent, render their shard through the C ABI, combine at a root and compare there with the single-GPU render:
tile shards bit for bit, sample shards (uneven counts included) against the count-weighted mean."""
import os
import sys

import numpy as np
import pytest

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, BOUNCES = 120, 72, 4


def _n_gpus():
    import torch
    return torch.cuda.device_count()


def _worker(rank, world, uid, out_dir):
    sys.path.insert(0, ROOT)
    import ray_tracer_b200 as rt
    from oracle import host_ref as hr
    ok = {}
    try:
        scene = hr.scene_room(np.float32(W) / np.float32(H))
        cam, sp, vs, ix, ms = scene

        def tracer(**shard):
            tr = rt.RayTracer(rank, W, H)
            tr.upload_scene(sp, vs, ix, ms); tr.set_camera(cam.view(rt.CAMERA_DT))
            tr.set_params(number_of_bounces=BOUNCES, rays_per_pixel=1, skybox=1, frames=0)
            tr.set_shard(**shard)
            return tr

        # --- tile shards, root 0 and root world-1; a second combine after more frames
        tr = tracer(tile_rank=rank, tile_world=world)
        tr.comm_init(world, rank, uid)
        tr.render_frames(0, 3)
        tr.combine(0)
        if rank == 0:
            one = tracer(); one.render_frames(0, 3); full = one.read_accum()
            ok["tile_root0"] = bool(np.array_equal(tr.read_combined().view(np.uint32), full.view(np.uint32)))
            ok["tile_accum_untouched"] = not np.array_equal(tr.read_accum(), full)
        tr.render_frames(3, 2)
        last = world - 1
        got8 = tr.combine_rgba8(last, is_root=(rank == last))
        tr.combine(last)
        if rank == last:
            one = tracer(); one.render_frames(0, 5); full = one.read_accum(); full8 = one.read_rgba8(); one.close()
            ok["tile_root_last"] = bool(np.array_equal(tr.read_combined().view(np.uint32), full.view(np.uint32)))
            ok["tile_rgba8"] = bool(np.array_equal(got8, full8))
        tr.sync()
        # --- sample shards with uneven counts (7 frames over `world` ranks), same communicator
        tr.set_shard(frame_rank=rank, frame_world=world)
        tr.clear()
        tr.render_frames(0, 7)
        np.save(os.path.join(out_dir, "part_%d.npy" % rank), tr.read_accum())
        np.save(os.path.join(out_dir, "count_%d.npy" % rank), np.array([tr.mean_count]))
        tr.combine(0)
        if rank == 0:
            np.save(os.path.join(out_dir, "sample_combined.npy"), tr.read_combined())
            o2 = tracer(); o2.render_frames(0, 7); np.save(os.path.join(out_dir, "sample_single.npy"), o2.read_accum()); o2.close()
        tr.sync()
        tr.comm_destroy()
        tr.close()
        ok["done"] = True
    except Exception as e:                                     # surfaced by the parent
        ok["error"] = repr(e)
    np.save(os.path.join(out_dir, "ok_%d.npy" % rank), np.array([repr(ok)]))


@pytest.mark.parametrize("world", [2, 4, 8])
def test_multi_process_combine_over_nccl(rtlib, tmp_path, world):
    if _n_gpus() < world:
        pytest.skip("needs %d GPUs" % world)
    import torch.multiprocessing as mp
    uid = rtlib.comm_unique_id()
    assert len(uid) == 128
    mp.spawn(_worker, args=(world, uid, str(tmp_path)), nprocs=world, join=True)
    oks = [eval(str(np.load(str(tmp_path / ("ok_%d.npy" % r)))[0])) for r in range(world)]
    for r, ok in enumerate(oks):
        assert "error" not in ok, (r, ok)
        assert ok.get("done"), (r, ok)
    assert oks[0]["tile_root0"] and oks[0]["tile_accum_untouched"]
    assert oks[world - 1]["tile_root_last"] and oks[world - 1]["tile_rgba8"]
    parts = [np.load(str(tmp_path / ("part_%d.npy" % r))) for r in range(world)]
    counts = [int(np.load(str(tmp_path / ("count_%d.npy" % r)))[0]) for r in range(world)]
    assert sum(counts) == 7 and counts == [len(range(r, 7, world)) for r in range(world)]
    want = None
    for p, n in zip(parts, counts):                            # rank order, separately rounded multiply and add
        if n == 0:
            continue
        term = (p * (np.float32(n) / np.float32(7))).astype(np.float32)
        want = term if want is None else (want + term).astype(np.float32)
    got = np.load(str(tmp_path / "sample_combined.npy"))
    assert np.array_equal(got.view(np.uint32), want.view(np.uint32))
    single = np.load(str(tmp_path / "sample_single.npy"))
    assert np.allclose(got, single, rtol=3e-6, atol=2e-6)     # == the sequential running mean up to f32 rounding


def test_comm_argument_errors(rtlib):
    L = rtlib._lib.lib()
    assert L.rt_comm_unique_id(None, 128) == rtlib._lib.RT_ERR_INVALID_ARGUMENT
    tr = rtlib.RayTracer(0, 16, 16)
    with pytest.raises(rtlib.RtError):
        tr.combine(0)                                          # no communicator
    with pytest.raises(rtlib.RtError):
        tr.comm_init(2, 5, b"\0" * 128)                        # rank out of range
    with pytest.raises(rtlib.RtError):
        tr.comm_init(1, 0, b"\0" * 12)                         # id size
    with pytest.raises(rtlib.RtError):
        tr.read_combined()                                     # nothing combined yet
    tr.close()


def test_single_rank_communicator(rtlib):
    """world = 1: the combine degenerates to a copy of the rank's own image (and its tone-mapped bytes)."""
    from oracle import host_ref as hr
    cam, sp, vs, ix, ms = hr.scene_balls(np.float32(W) / np.float32(H))
    tr = rtlib.RayTracer(0, W, H)
    tr.upload_scene(sp, vs, ix, ms); tr.set_camera(cam.view(rtlib.CAMERA_DT))
    tr.set_params(number_of_bounces=3, rays_per_pixel=1, skybox=1, frames=0)
    tr.comm_init(1, 0, rtlib.comm_unique_id())
    tr.render_frames(0, 2)
    tr.combine(0)
    assert np.array_equal(tr.read_combined().view(np.uint32), tr.read_accum().view(np.uint32))
    assert np.array_equal(tr.combine_rgba8(0), tr.read_rgba8())
    tr.comm_destroy(); tr.close()
