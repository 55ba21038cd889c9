"""Multi-GPU sharding of a frame: the samples of every pixel are split across ranks.

One process per GPU; each rank renders the global sample range sample_split() gives it into its own fp32
radiance buffer; ONE collective per frame — a sum-reduce of that buffer to rank 0 over NCCL (NVLink /
NVSwitch) — then rank 0 resolves.  Philox is keyed by the GLOBAL sample index, so the frame does not
depend on the number of ranks (up to fp32 summation order).  There is no data-path collective inside
the render itself: paths are independent.
"""


def sample_split(spp, world_size, rank):
    """-> (sample_begin, sample_count) of `rank`: spp/world each, the remainder to the low ranks."""
    if spp < 0 or world_size < 1 or not (0 <= rank < world_size):
        raise ValueError("bad sample split %r" % ((spp, world_size, rank),))
    base, rem = divmod(spp, world_size)
    count = base + (1 if rank < rem else 0)
    begin = rank * base + min(rank, rem)
    return begin, count


def reduce_frame(accum, dst=0):
    """Sum-reduce the per-rank radiance buffers to `dst` (in place there).  No-op without a process group."""
    import torch.distributed as dist
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.reduce(accum, dst=dst, op=dist.ReduceOp.SUM)
    return accum
