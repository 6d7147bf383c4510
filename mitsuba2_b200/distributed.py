"""Multi-GPU sharding of the path: one process per GPU, every rank holds the whole scene, the sample
indices of every pixel are split over the ranks, and the XYZAW film sums are all-reduced once per render (NCCL on
GPUs; the host logic is backend-agnostic so the world_size-2 CPU tests run it over gloo).

The reference has no multi-device path; the closest counterpart is the block loop of
`SamplingIntegrator::render` (src/librender/integrator.cpp:52-173) whose per-block `Film::put` merge this replaces.
"""
from __future__ import annotations

import os


def world_info(group=None):
    """(rank, world) from torch.distributed when initialised, else from the torchrun environment, else (0, 1)."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(group), dist.get_world_size(group)
    except ImportError:
        pass
    return int(os.environ.get("RANK", "0")), int(os.environ.get("WORLD_SIZE", "1"))


def sample_share(spp_total: int, rank: int, world: int):
    """Contiguous share [begin, begin+count) of the per-pixel sample indices [0, spp_total) owned by `rank`.
    Shares are disjoint, cover the range, and differ in size by at most one."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    base, rem = divmod(int(spp_total), world)
    begin = rank * base + min(rank, rem)
    return begin, base + (1 if rank < rem else 0)


def render_sharded(render_xyzaw, spp_total: int, seed: int = 0, group=None):
    """Render this rank's share and sum the films over the ranks.

    `render_xyzaw(spp_begin, spp_count, spp_total, seed)` must return a flat torch tensor of XYZAW sums (on the
    device the process group communicates on).  Every rank returns the reduced film; because the film holds sums
    and the weight channel, reduce-then-develop equals the single-device render up to float association."""
    import torch.distributed as dist
    rank, world = world_info(group)
    begin, count = sample_share(spp_total, rank, world)
    film = render_xyzaw(begin, count, spp_total, seed)
    if world > 1:
        dist.all_reduce(film, op=dist.ReduceOp.SUM, group=group)
    return film


def allreduce_gradients(grads, group=None):
    """Sum gradient tensors (texture / vertex buffers of mtsb_render_grad) over the ranks in place; every rank then
    applies the identical optimiser update, so parameters never need to be broadcast."""
    import torch.distributed as dist
    _, world = world_info(group)
    if world > 1:
        for g in grads:
            dist.all_reduce(g, op=dist.ReduceOp.SUM, group=group)
    return grads
