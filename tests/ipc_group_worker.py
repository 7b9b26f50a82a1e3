"""torchrun worker of tests/test_multi_gpu_gpu.py::test_ipc_group_under_torchrun: one process per GPU, CUDA IPC peers.
Rank 0 compares the group's frames with (a) the per-rank accumulators added in rank order (bit for bit) and (b) a
single-context render of all samples (1e-6)."""
import os
import sys

import numpy as np
import torch
import torch.distributed as dist

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
from sparkle_b200 import _abi as A  # noqa: E402
from sparkle_b200 import multi_gpu, scenes  # noqa: E402
from sparkle_b200.device import CudaContext  # noqa: E402
from sparkle_b200.scene import orbit_camera  # noqa: E402

rank, world, local = int(os.environ["RANK"]), int(os.environ["WORLD_SIZE"]), int(os.environ["LOCAL_RANK"])
torch.cuda.set_device(local)
dist.init_process_group("nccl", device_id=torch.device("cuda", local))
w, h, spp = 640, 360, 9
sb, camkw = scenes.testscene()
desc = sb.finalize()
cam = orbit_camera(w, h, **camkw)
ctx = CudaContext(w, h, device=local)
ctx.upload_scene(desc)
ctx.set_camera(cam)
group = multi_gpu.ipc_group(ctx, rank, world)
first, count, stride = multi_gpu.sample_slice(spp, rank, world)
frames = []
for frame in range(3):  # several frames: the arrive / depart epochs advance, accumulators are reused
    ctx.reset_accumulation()
    ctx.render(first, count, sample_stride=stride, max_bounce=8)
    part = torch.from_numpy(ctx.resolve(A.RESOLVE_LINEAR_F32, total_samples=1)).cuda()
    parts = [torch.empty_like(part) for _ in range(world)]
    dist.all_gather(parts, part)
    lin = group.resolve(A.RESOLVE_LINEAR_F32, spp)
    f16 = group.resolve(A.RESOLVE_TONEMAP_F16, spp)
    if rank == 0:
        expect = parts[0].cpu().numpy()
        for p in parts[1:]:
            expect = expect + p.cpu().numpy()
        expect = expect / np.float32(spp)
        assert np.array_equal(lin.view(np.uint32), expect.view(np.uint32)), "reduce order"
        frames.append((lin, f16))
if rank == 0:
    one = CudaContext(w, h, device=local)
    one.upload_scene(desc)
    one.set_camera(cam)
    one.render(0, spp, max_bounce=8)
    single = one.resolve(A.RESOLVE_LINEAR_F32)
    assert np.allclose(frames[0][0], single, rtol=1e-6, atol=1e-6), float(np.abs(frames[0][0] - single).max())
    for lin, f16 in frames[1:]:
        assert np.array_equal(lin, frames[0][0]) and np.array_equal(f16, frames[0][1])
    one.close()
    print("ipc group ok", flush=True)
dist.barrier()
group.close()
ctx.close()
dist.destroy_process_group()
