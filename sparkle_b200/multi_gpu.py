"""Multi-GPU partitioning of the path. One process per GPU; the scene is replicated; sample s of pixel p
depends only on (p, s) through its seed (CPURenderer.cpp:442-443), so ranks take interleaved sample indices and the
accumulators are summed with ONE reduce to rank 0. Ray batches split into contiguous slices with no collective.
torch.distributed is plumbing only (NCCL on GPUs, gloo in the CPU tests)."""


def sample_slice(total_spp, rank, world):
    """(first_sample, sample_count, sample_stride) of this rank: samples {rank, rank + world, ...} below total_spp.
    Uneven totals are allowed: the first total_spp % world ranks take one more sample."""
    if not (0 <= rank < world) or total_spp < 0:
        raise ValueError("bad rank/world/spp")
    count = (total_spp - rank + world - 1) // world if total_spp > rank else 0
    return rank, count, world


def ray_slice(num_rays, rank, world):
    """[begin, end) of this rank's contiguous share of a ray batch"""
    base, extra = divmod(num_rays, world)
    begin = rank * base + min(rank, extra)
    return begin, begin + base + (1 if rank < extra else 0)


def reduce_accumulator(accum_tensor, world, dst=0):
    """sum of the per-rank accumulators on `dst` (ncclReduce over NVLink on GPUs). In place; no-op for one rank."""
    if world > 1:
        import torch.distributed as dist
        dist.reduce(accum_tensor, dst=dst, op=dist.ReduceOp.SUM)
    return accum_tensor
