"""Multi-GPU partition of the path (no reference counterpart: upstream is single-device).

Every pixel and every dispatch index is independent (per-pixel RNG seeded from pixel and
frames, wgsl:308), so the work shards without any data-path collective; the only exchange
is the combine of the per-GPU accumulation images when an image is wanted:

  tile mode   — 8x8-pixel tiles dealt round-robin (tile t -> rank t % world, as the kernel
                does for rt_set_shard); images are disjoint, the combine is a plain sum and
                is bit-exact with the single-GPU image;
  sample mode — dispatch f belongs to rank f % world and is blended with weight
                1/(f // world + 1), i.e. every rank keeps the running mean of its own
                dispatches; the combine is the count-weighted mean of the rank means (equal
                to the reference's sequential running mean up to f32 rounding).

The combine runs over torch.distributed (NCCL over NVLink on the GPU box, gloo in the CPU
tests); tensors may live on either device.
"""
import numpy as np

TILE = 8


def tile_owner(width, height, world):
    """(H, W) int32: rank owning each pixel under tile sharding."""
    tiles_x = (width + TILE - 1) // TILE
    ty, tx = np.meshgrid(np.arange(height) // TILE, np.arange(width) // TILE, indexing="ij")
    return ((ty * tiles_x + tx) % world).astype(np.int32)


def frames_of_rank(first_frame, n_frames, rank, world):
    """Dispatch indices of [first_frame, first_frame + n_frames) rendered by `rank` in sample mode."""
    return [f for f in range(first_frame, first_frame + n_frames) if f % world == rank]


def combine_tile_shards(accum, group=None):
    """Sum of disjoint per-rank images, in place on every rank."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(accum, op=dist.ReduceOp.SUM, group=group)
    return accum


def combine_sample_shards(accum, n_own, n_total, group=None):
    """Count-weighted mean of the per-rank running means, in place on every rank."""
    import torch.distributed as dist
    if dist.is_initialized() and dist.get_world_size(group) > 1:
        accum.mul_(float(n_own) / float(n_total))
        dist.all_reduce(accum, op=dist.ReduceOp.SUM, group=group)
    return accum
