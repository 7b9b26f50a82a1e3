"""A/B of SPK_OPT_TRACE_SMEM_TOP (the top levels of the hottest tree staged in shared memory by the ray-batch kernel):
soup-1M incoherent rays (host SAH tree and device LBVH), terrain-1M coherent rays, city-10M incoherent rays. Best of 3 device
times per setting; hits of every setting are compared bit for bit with the first."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from sparkle_b200 import _abi as A, geometry, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import SceneBuilder

N = int(os.environ.get("SPK_RAYS", 16 << 20))
SETTINGS = [int(x) for x in os.environ.get("SPK_TOPS", "0,64,128,256,512").split(",")]


def run(name, sb, rays):
    ctx = CudaContext(64, 64)
    ctx.upload_scene(sb.finalize())
    d_rays = torch.from_numpy(rays).cuda()
    hits = torch.empty((len(rays), 5), dtype=torch.float32, device="cuda")
    first = None
    for top in SETTINGS:
        ctx.set_option(A.OPT_TRACE_SMEM_TOP, top)
        best = 1e30
        for rep in range(4):
            ctx.trace_closest_device(d_rays.data_ptr(), len(rays), hits.data_ptr(), 0)
            ctx.synchronize()
            if rep:
                best = min(best, ctx.stats()["last_ms"])
        h = hits.cpu().numpy().view(np.uint32)
        same = "" if first is None else (" same bits" if np.array_equal(h, first) else " HITS DIFFER")
        first = h if first is None else first
        print(f"{name}: top {top:4d} pairs: {best:8.3f} ms = {len(rays) / best / 1e3:7.0f} Mrays/s{same}", flush=True)
    ctx.close()


rays = geometry.random_rays(N, seed=2)
sb, _ = scenes.soup(1_000_000)
run("soup-1M host SAH tree, incoherent", sb, rays)
dev = SceneBuilder(device_build=True)
dev.add_mesh_instance(dev.add_mesh(sb.meshes[0]), dev.lambertian())
run("soup-1M device LBVH, incoherent", dev, rays)
sb, _ = scenes.city()
crays = geometry.random_rays(N, seed=3, extent=40.0)
crays[:, 2] = np.abs(crays[:, 2]) * 0.3 + 0.5
run("city-10M (TLAS staged), incoherent", sb, crays)
