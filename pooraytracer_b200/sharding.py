"""Multi-GPU plumbing of the render path: the scene is replicated, the samples of
every pixel are sharded across ranks, and the per-rank float64 accumulators are summed by ONE
collective. There is no exchange inside the path loop, so this is all the multi-GPU logic there is.

The counter-based generator is keyed by (pixel, absolute sample index): rank r rendering
[begin, end) produces exactly the samples a single GPU would have produced for those indices, so
the reduced image equals the single-GPU image up to the order of the final FP64 additions.
"""
from __future__ import annotations


def sample_range(step: int, spp_per_step: int, rank: int, world: int) -> tuple[int, int]:
    """Weak scaling: every rank renders `spp_per_step` samples per step; step s of the whole job covers
    the absolute sample indices [s*S*world, (s+1)*S*world), rank r takes the r-th slice of it."""
    begin = step * spp_per_step * world + rank * spp_per_step
    return begin, begin + spp_per_step


def split_samples(total_spp: int, rank: int, world: int) -> tuple[int, int]:
    """Strong scaling: the r-th of `world` contiguous, balanced slices of [0, total_spp)
    (sizes differ by at most one: 500 spp over 8 GPUs = 63,63,63,63,62,62,62,62)."""
    base, extra = divmod(total_spp, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def ray_range(n_rays: int, rank: int, world: int) -> tuple[int, int]:
    """Ray queries: the r-th of `world` contiguous, balanced ranges of a batch of `n_rays`;
    no collective, every rank returns the results of its own range."""
    return split_samples(n_rays, rank, world)


def reduce_image(acc, dst: int = 0):
    """The single collective: sum of the accumulators on rank `dst` (NCCL over NVLink on GPUs, gloo in
    the CPU tests). `acc` is a torch tensor; no-op when torch.distributed is not initialised."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(acc, dst=dst, op=dist.ReduceOp.SUM)
    return acc
