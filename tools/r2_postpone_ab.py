"""A/B of builds of libsparkle_cuda (SPK_CUDA_LIB): the wavefront tracer on city-10M at 4K x 16 spp, ray batches (16 Mi
rays, device buffers) on soup-1M (incoherent), city-10M (incoherent) and terrain-1M (camera grid). Prints ms and a hash of
the hits / image, so that runs of different builds can be compared line by line."""
import hashlib, os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

lib = os.environ.get("SPK_CUDA_LIB", "default")
what = sys.argv[1].split(",") if len(sys.argv) > 1 else ["city_wf", "soup", "city", "terrain"]

def h(a):
    return hashlib.sha1(np.ascontiguousarray(a).tobytes()).hexdigest()[:12]

def trace(name, sb, rays):
    ctx = CudaContext(64, 64); ctx.set_stream(torch.cuda.current_stream().cuda_stream); ctx.upload_scene(sb.finalize())
    n = rays.shape[0]
    hits = torch.empty((n, 5), dtype=torch.float32, device="cuda")
    for flags, tag in ((0, "culled"), (A.FLAG_EXACT_TRAVERSAL, "exact")):
        best = 1e30
        for rep in range(4):
            ctx.trace_closest_device(rays.data_ptr(), n, hits.data_ptr(), flags); ctx.synchronize()
            if rep: best = min(best, ctx.stats()["last_ms"])
        print(f"{lib}: trace {name} {tag}: {best:.2f} ms = {n/best/1e3:.0f} Mrays/s hits {h(hits.cpu().numpy())}", flush=True)
    ctx.close()

def rand_rays(n, extent, seed):
    g = torch.Generator(device="cuda").manual_seed(seed)
    rays = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
    rays[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * (2 * extent) - extent
    d = torch.randn((n, 3), generator=g, device="cuda")
    rays[:, 4:7] = d / d.norm(dim=1, keepdim=True)
    return rays

N = 16 << 20
if "city_wf" in what or "city" in what:
    city, camkw = scenes.city()
if "city_wf" in what:
    w, hh, spp = 3840, 2160, 16
    ctx = CudaContext(w, hh); ctx.upload_scene(city.finalize()); ctx.set_camera(orbit_camera(w, hh, **camkw))
    best = 1e30
    for rep in range(3):
        ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=8, flags=A.FLAG_WAVEFRONT); ctx.synchronize()
        if rep: best = min(best, ctx.stats()["last_ms"])
    img = ctx.resolve(A.RESOLVE_LINEAR_F32)
    print(f"{lib}: wavefront city-10M 4K {spp} spp: {best:.2f} ms = {w*hh*spp/best/1e3:.1f} Mpaths/s mean {float(img[..., :3].mean()):.9f}", flush=True)
    ctx.close()
if "soup" in what:
    sb, _ = scenes.soup(1_000_000)
    trace("soup-1M incoherent", sb, rand_rays(N, 1.5, 2))
if "city" in what:
    r = rand_rays(N, 30.0, 4); r[:, 2] = r[:, 2].abs() * 0.3 + 0.1
    trace("city-10M incoherent", city, r)
if "terrain" in what:
    sb, camkw = scenes.terrain(1000, 500)
    sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
    import bench
    rays = bench.camera_grid_rays(torch, orbit_camera(4096, 4096, **camkw), 4096, torch.device("cuda"))
    trace("terrain-1M camera grid", sb, rays.contiguous())
