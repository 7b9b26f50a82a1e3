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


class Group:
    """SpkGroup of include/spk_cuda.h: the accumulators of all ranks are summed, tone-mapped and gathered at the root by one
    kernel per GPU over NVLink peer memory (csrc/spk_group.cu). torch.distributed / any transport only carries the
    192-byte peer handles once, at set-up."""

    def __init__(self, handle, lib, root_ctx, is_root):
        self.handle, self.lib, self.ctx, self.is_root = handle, lib, root_ctx, is_root

    @classmethod
    def local(cls, ctxs):
        """all contexts in this process (one per GPU; the first is the root)"""
        import ctypes as C
        from . import _abi as A
        lib = A.cuda_lib()
        arr = (A.CtxP * len(ctxs))(*[c.handle for c in ctxs])
        g = A.GroupP()
        A.check(lib.spk_cuda_group_create_local(arr, len(ctxs), C.byref(g)))
        return cls(g, lib, ctxs[0], True)

    @staticmethod
    def export(ctx):
        """this rank's peer handle as bytes (to be all-gathered by the host)"""
        import ctypes as C
        from . import _abi as A
        h = A.PeerHandle()
        A.check(ctx.lib.spk_cuda_group_export(ctx.handle, C.byref(h)))
        return bytes(h.bytes)

    @classmethod
    def ipc(cls, ctx, rank, world, handles):
        """one process per GPU: `handles` = the exported handles of ranks 0..world-1 in rank order"""
        import ctypes as C
        from . import _abi as A
        arr = (A.PeerHandle * world)()
        for r, b in enumerate(handles):
            C.memmove(C.addressof(arr[r]), bytes(b), A.PEER_HANDLE_BYTES)
        g = A.GroupP()
        A.check(ctx.lib.spk_cuda_group_create(ctx.handle, rank, world, arr, C.byref(g)))
        return cls(g, ctx.lib, ctx, rank == 0)

    def resolve(self, fmt, total_samples, out=None, readback=True):
        """collective: every rank calls it; the root returns the frame, the others None. readback=False leaves the frame
        in the root's device buffer and returns without blocking."""
        import numpy as np
        from . import _abi as A
        if self.is_root and readback and out is None:
            out = np.empty((self.ctx.height, self.ctx.width, 4), self.ctx.RESOLVE_DTYPES[fmt])
        ptr, n = (out.ctypes.data, out.nbytes) if (self.is_root and readback) else (None, 0)
        A.check(self.lib.spk_cuda_group_resolve(self.handle, fmt, total_samples, ptr, n))
        return out if (self.is_root and readback) else None

    def set_async(self, enabled=True):
        from . import _abi as A
        A.check(self.lib.spk_cuda_group_set_async(self.handle, int(enabled)))

    def close(self):
        if self.handle:
            self.lib.spk_cuda_group_destroy(self.handle)
            self.handle = None


def ipc_group(ctx, rank, world):
    """set-up under torchrun: all-gather the peer handles over torch.distributed, then map the peers"""
    import torch
    import torch.distributed as dist
    mine = torch.frombuffer(bytearray(Group.export(ctx)), dtype=torch.uint8).cuda()
    everyone = [torch.empty_like(mine) for _ in range(world)]
    dist.all_gather(everyone, mine)
    return Group.ipc(ctx, rank, world, [bytes(t.cpu().numpy().tobytes()) for t in everyone])
