"""A/B of megakernel forms incl. 4-warp CTAs (1048 = 4 warps x 8 CTAs/SM, 1046; built by adding SPK_SORTED_CASE(4, 8, 0) and
(4, 6, 0) to spk_render.cu) on TestScene and room. Measured (B200, 1080p): TestScene 64 spp 1084 32.97 / 1048 36.73 / 1046 39.35 /
1162 33.20 ms; room 32 spp 119.8 / 120.1 / 127.9 / 125.2 ms: smaller CTAs wait less at the barrier but sort worse; not shipped."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
import numpy as np
for name, (w, h, b, spp) in {"testscene": (1920, 1080, 8, 64), "room": (1920, 1080, 32, 32)}.items():
    sb, camkw = getattr(scenes, name)()
    ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **camkw))
    ref = None
    for form in (1084, 1048, 1046, 1162):
        ctx.set_option(A.OPT_MEGAKERNEL_FORM, form)
        best = 1e30
        for rep in range(6):
            ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=b); ctx.synchronize()
            if rep > 1: best = min(best, ctx.stats()["last_ms"])
        img = ctx.resolve(A.RESOLVE_LINEAR_F32)
        if ref is None: ref = img
        print(f"{name} {spp} spp form {form}: {best:.3f} ms = {w*h*spp/best/1e3:.0f} Mpaths/s {'same bits' if np.array_equal(img, ref) else 'DIFFERS'}", flush=True)
    ctx.close()
