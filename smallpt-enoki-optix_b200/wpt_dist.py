"""Multi-GPU plumbing: samples-per-pixel shard across ranks, one rank per GPU.
Work unit = (pixel, global spp index); paths never interact and the film is a sum, so the only
exchange step is ONE reduce of the partial films (3*W*H floats) onto rank 0.  The scene/BVH is
replicated.  torch.distributed is plumbing: NCCL over NVLink on the GPU box, gloo in the CPU tests.
The reference has no counterpart (single process, single GPU)."""


def spp_range_strong(rank, world, spp_total):
    """Strong scaling: split [0, spp_total) into `world` contiguous, disjoint ranges that differ in
    size by at most one.  Returns (spp_begin, spp_count)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(spp_total, world)
    begin = rank * base + min(rank, rem)
    return begin, base + (1 if rank < rem else 0)


def spp_range_weak(rank, world, spp_per_rank):
    """Weak scaling (bench.py): every rank renders spp_per_rank samples of the same estimator;
    rank r owns global indices [r*spp_per_rank, (r+1)*spp_per_rank)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return rank * spp_per_rank, spp_per_rank


def reduce_film(film, dst=0):
    """Sum the partial radiance buffers onto `dst` (ncclReduce / gloo reduce). No-op when not
    initialised or world == 1.  Returns the total number of ranks."""
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return 1
    dist.reduce(film, dst=dst, op=dist.ReduceOp.SUM)
    return dist.get_world_size()
