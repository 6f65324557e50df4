"""N > 1 host logic on CPU: two gloo ranks shard the work exactly as the GPU ranks do
(ray_tracer_b200/sharding.py), render their share with the oracle, combine over
torch.distributed, and must reproduce the single-process image."""
import os
import socket
import sys

import numpy as np
import pytest

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
W, H, FRAMES, BOUNCES = 40, 24, 4, 3


def _frag(ro, scene, f):
    """frag of dispatch f alone: frames = -f overwrites (wgsl:64-66) and seeds with abs(frames) = f."""
    return ro.render(scene, W, H, 1, BOUNCES, 1, 1, -f)


def _worker(rank, world, port, mode, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    from oracle import host_ref as hr, rt_oracle as ro
    from ray_tracer_b200 import sharding
    dist.init_process_group("gloo", init_method="tcp://127.0.0.1:%d" % port, rank=rank, world_size=world)
    scene = hr.scene_room(np.float32(W) / np.float32(H))
    if mode == "tile":
        full = ro.render(scene, W, H, FRAMES, BOUNCES, 1, 1, 0, nthreads=2)
        mine = np.where((sharding.tile_owner(W, H, world) == rank)[..., None], full, np.float32(0)).astype(np.float32)
        t = torch.from_numpy(mine)
        sharding.combine_tile_shards(t)
    else:
        own = sharding.frames_of_rank(0, FRAMES, rank, world)
        acc = np.zeros((H, W, 4), np.float32)
        for f in own:
            k = f // world
            fr = _frag(ro, scene, f)
            if k >= 1:
                w = np.float32(1.0) / np.float32(k + 1)
                acc = (acc * (np.float32(1.0) - w) + fr * w).astype(np.float32)
            else:
                acc = fr
        t = torch.from_numpy(acc.copy())
        sharding.combine_sample_shards(t, len(own), FRAMES)
    np.save(os.path.join(out_dir, "%s_%d.npy" % (mode, rank)), t.numpy())
    dist.destroy_process_group()


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    p = s.getsockname()[1]
    s.close()
    return p


@pytest.mark.parametrize("mode", ["tile", "sample"])
def test_two_rank_shard_and_combine(oracle, tmp_path, mode):
    import torch.multiprocessing as mp
    from oracle import host_ref as hr
    world = 2
    mp.spawn(_worker, args=(world, _free_port(), mode, str(tmp_path)), nprocs=world, join=True)
    ref = oracle.render(hr.scene_room(np.float32(W) / np.float32(H)), W, H, FRAMES, BOUNCES, 1, 1, 0)
    outs = [np.load(str(tmp_path / ("%s_%d.npy" % (mode, r)))) for r in range(world)]
    assert np.array_equal(outs[0], outs[1])
    if mode == "tile":
        assert np.array_equal(outs[0], ref)                      # disjoint tiles: bit-exact
    else:
        # rank means vs the sequential running mean: equal up to f32 rounding
        assert np.allclose(outs[0], ref, rtol=2e-6, atol=1e-6)


def test_tile_owner_and_frame_split():
    from ray_tracer_b200 import sharding
    o = sharding.tile_owner(20, 12, 3)
    assert o.shape == (12, 20) and o[0, 0] == 0 and o[0, 8] == 1 and o[0, 16] == 2 and o[8, 0] == 0 and o[8, 8] == 1
    assert (o[:8, :8] == 0).all()
    assert sharding.frames_of_rank(0, 8, 1, 4) == [1, 5] and sharding.frames_of_rank(3, 4, 0, 2) == [4, 6]
    allf = sorted(sum((sharding.frames_of_rank(5, 11, r, 3) for r in range(3)), []))
    assert allf == list(range(5, 16))
