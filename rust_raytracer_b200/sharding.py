"""Multi-GPU sharding of capture_image: by SAMPLE INDEX (north_star).

Rank g of G renders the global samples [begin, begin+count) of every pixel; the Philox key carries the global
sample index, so the union over ranks is the single-GPU image up to float summation order.  The only data-path
exchange is one sum-reduce of the float accumulation buffer (NCCL over NVLink; gloo in the CPU tests).
"""


def spp_range(rank, world_size, spp_total, spp_begin=0):
    """Contiguous, disjoint, exhaustive split of [spp_begin, spp_begin + spp_total) — remainder to the low ranks."""
    if world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad rank/world_size")
    if spp_total < 0:
        raise ValueError("negative spp")
    base, rem = divmod(spp_total, world_size)
    count = base + (1 if rank < rem else 0)
    begin = spp_begin + rank * base + min(rank, rem)
    return begin, count


def reduce_accum(accum, dst=0, group=None):
    """Sum the (H, W, 4) accumulation tensors of all ranks onto `dst` (torch.distributed; in place)."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM, group=group)
    return accum
