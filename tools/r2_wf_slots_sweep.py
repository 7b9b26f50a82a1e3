"""Sweep of the wavefront tracer's slot pool (SPK_OPT_WAVEFRONT_SLOTS) on city-10M at 4K x 32 spp."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
w, h, spp = 3840, 2160, 32
sb, camkw = scenes.city()
ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **camkw))
for mi in (4, 8, 16, 32, 64):
    ctx.set_option(A.OPT_WAVEFRONT_SLOTS, mi << 20)
    best = 1e30
    for rep in range(3):
        ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=8, flags=A.FLAG_WAVEFRONT); ctx.synchronize()
        if rep: best = min(best, ctx.stats()["last_ms"])
    print(f"{mi} Mi slots: {best:.2f} ms = {w*h*spp/best/1e3:.1f} Mpaths/s", flush=True)
