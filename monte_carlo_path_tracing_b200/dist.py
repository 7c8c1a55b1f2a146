"""Multi-GPU: shard samples-per-pixel across ranks, one NCCL reduce of the float film at the end.

Every (pixel, sample) is independent in the reference (main.cpp:567-577: samples only meet in
`sum = sum + I`), so rank r of G renders ALL pixels for the global sample indices
[offset_r, offset_r + n_r) -- the counter-based RNG is keyed on the global index, so the result does not
depend on G beyond FP32 summation order -- normalised by the GLOBAL spp, and the W*H*3 float buffers are
summed onto rank 0 with a single `torch.distributed.reduce` (NCCL over NVLink on GPUs, gloo on CPU tests).
The scene is replicated per GPU.  There is no data-path collective inside the render.
"""
from __future__ import annotations

import numpy as np


def shard_spp(total_spp: int, world_size: int, rank: int):
    """(offset, count) of the contiguous global sample range owned by `rank` (remainder to low ranks)."""
    base, rem = divmod(int(total_spp), int(world_size))
    count = base + (1 if rank < rem else 0)
    offset = rank * base + min(rank, rem)
    return offset, count


def reduce_film(film, dst: int = 0, group=None):
    """Sum the per-rank partial films onto `dst` (in place).  `film`: torch tensor (CUDA -> NCCL, CPU -> gloo)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return film


def render_sharded(scene, cam, strategy: int, total_spp: int, rank: int, world_size: int, film, stream: int = 0,
                   spp_base: int = 0, **kw):
    """Render this rank's share of `total_spp` into `film` (torch CUDA float32 tensor of W*H*3) and reduce.

    Returns the rank's McptStats dict.  After the call rank 0's `film` holds the full image."""
    off, cnt = shard_spp(total_spp, world_size, rank)
    if cnt > 0:
        st = scene.render_device(cam, strategy, cnt, film.data_ptr(), stream=stream, spp_offset=spp_base + off,
                                 spp_total=total_spp, **kw)
    else:
        film.zero_()
        st = {}
    reduce_film(film, 0)
    return st


def combine_host(partials):
    """CPU reference of the reduce: sum of per-rank partial films (numpy)."""
    out = np.zeros_like(partials[0], dtype=np.float64)
    for p in partials:
        out += p
    return out
