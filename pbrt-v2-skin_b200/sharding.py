"""Multi-GPU partitioning of the hot path: contiguous equal shards of the irradiance
points and of the camera samples, and the in-place all-gather of the irradiance set.  Pure
host logic over torch.distributed; used by bench.py (NCCL) and tested on CPU with gloo."""


def shard_range(n, rank, world):
    """Contiguous shard [first, first+count) of n items; every shard but the last has ceil(n/world)."""
    per = (n + world - 1) // world
    first = min(rank * per, n)
    return first, min(per, n - first), per


def tile_round_robin(n, tile, rank, world):
    """Indices of the camera samples of `rank` when tiles of `tile` consecutive samples are dealt round-robin
    (the reference's own tile = one SamplerRendererTask window).  Contiguous halves of the image carry
    very different amounts of skin; interleaved tiles balance the Mo work across ranks."""
    import numpy as np
    tiles = (n + tile - 1) // tile
    mine = np.arange(rank, tiles, world, dtype=np.int64)
    idx = (mine[:, None] * tile + np.arange(tile, dtype=np.int64)[None, :]).reshape(-1)
    return idx[idx < n]


def allgather_inplace(dist, full, rank, per):
    """full: tensor of world*per rows; rows [rank*per, (rank+1)*per) hold this rank's shard.
    After the call every rank holds every shard (NCCL/gloo all_gather_into_tensor, no staging copy)."""
    dist.all_gather_into_tensor(full, full[rank * per:(rank + 1) * per])
    return full


def allgather_inplace_async(dist, full, rank, per):
    """Same, returning the work handle: the caller overlaps the transfer with work that does not read `full`
    (bench.py: octree topology from the positions while the 8x larger irradiance rows are in flight)."""
    return dist.all_gather_into_tensor(full, full[rank * per:(rank + 1) * per], async_op=True)
