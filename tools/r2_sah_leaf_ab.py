"""A/B: largest leaf of the host SAH builder (SceneBuilder(max_leaf=...)) on the bench workloads: C2 TestScene and C5 room
(megakernel 8x4), C4 city-10M (wavefront, 4K), C3 soup-1M (16 Mi incoherent rays)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

leaves = [int(x) for x in sys.argv[1].split(",")] if len(sys.argv) > 1 else [1, 2, 3, 4, 6]
what = sys.argv[2].split(",") if len(sys.argv) > 2 else ["testscene", "room", "city", "soup"]
for leaf in leaves:
    for name in what:
        if name == "soup":
            sb, _ = scenes.soup(1_000_000, max_leaf=leaf)
            ctx = CudaContext(64, 64); ctx.set_stream(torch.cuda.current_stream().cuda_stream); ctx.upload_scene(sb.finalize())
            n = 16 << 20
            g = torch.Generator(device="cuda").manual_seed(2)
            rays = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
            rays[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * 3.0 - 1.5
            d = torch.randn((n, 3), generator=g, device="cuda"); rays[:, 4:7] = d / d.norm(dim=1, keepdim=True)
            hits = torch.empty((n, 5), dtype=torch.float32, device="cuda")
            best = 1e30
            for rep in range(4):
                ctx.trace_closest_device(rays.data_ptr(), n, hits.data_ptr(), 0); ctx.synchronize()
                if rep: best = min(best, ctx.stats()["last_ms"])
            print(f"max_leaf {leaf}: soup-1M 16 Mi rays {best:.2f} ms = {n/best/1e3:.0f} Mrays/s", flush=True)
            ctx.close(); del rays, hits
            continue
        w, h, b, spp, flags = {"testscene": (1920, 1080, 8, 32, 0), "room": (1920, 1080, 32, 16, 0),
                               "city": (3840, 2160, 8, 16, A.FLAG_WAVEFRONT)}[name]
        sb, camkw = getattr(scenes, name)(max_leaf=leaf)
        ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **camkw))
        ctx.set_option(A.OPT_MEGAKERNEL_FORM, 1084)
        best = 1e30
        for rep in range(4):
            ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=b, flags=flags); ctx.synchronize()
            if rep: best = min(best, ctx.stats()["last_ms"])
        print(f"max_leaf {leaf}: {name} {w}x{h} {spp} spp {best:.2f} ms = {w*h*spp/best/1e3:.1f} Mpaths/s", flush=True)
        ctx.close()
