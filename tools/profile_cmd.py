"""Short command for `ncu --set full`: one render launch (TestScene 1920x1080, 16 spp) and one ray-batch launch
(soup-1M, 8M incoherent rays, device-resident), each after a warm-up launch. Prints device times from CUDA events."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from sparkle_b200 import _abi as A, geometry, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

what = sys.argv[1] if len(sys.argv) > 1 else "both"
if what in ("both", "render"):
    w, h = 1920, 1080
    sb, camkw = scenes.testscene()
    ctx = CudaContext(w, h)
    ctx.upload_scene(sb.finalize())
    ctx.set_camera(orbit_camera(w, h, **camkw))
    for k in range(2):
        ctx.reset_accumulation()
        ctx.render(0, int(os.environ.get("SPK_SPP", 16)), max_bounce=8)
        ctx.synchronize()
        print("render spp ms", ctx.stats()["last_ms"])
    ctx.close()
if what in ("both", "trace"):
    n_tris = int(os.environ.get("SPK_TRIS", 1_000_000))
    n_rays = int(os.environ.get("SPK_RAYS", 8 << 20))
    sb, _ = scenes.soup(n_tris)
    ctx = CudaContext(64, 64)
    ctx.upload_scene(sb.finalize())
    rays = torch.from_numpy(geometry.random_rays(n_rays, seed=2)).cuda()
    hits = torch.empty((n_rays, 5), dtype=torch.float32, device="cuda")
    torch.cuda.synchronize()
    for k in range(2):
        ctx.trace_closest_device(rays.data_ptr(), n_rays, hits.data_ptr(), A.FLAG_THREAD_PER_RAY if os.environ.get('SPK_TPR') else 0)
        ctx.synchronize()
        ms = ctx.stats()["last_ms"]
        print("trace", n_rays, "rays ms", ms, "Mrays/s", n_rays / ms / 1e3)
    ctx.reset_stats()
    ctx.trace_closest_device(rays.data_ptr(), n_rays, hits.data_ptr(), A.FLAG_COUNTERS)
    st = ctx.stats()
    nv, tt = st["node_visits"] / n_rays, st["tri_tests"] / n_rays
    print("node visits/ray", nv, "tri tests/ray", tt, "alg bytes/ray", 32 + 20 + 64 * nv + 48 * tt + 128)
    ctx.close()
