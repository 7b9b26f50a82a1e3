"""Wavefront vs megakernel on room (C5 shape: 1080p, depth 32) at 8 / 32 spp."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
w, h = 1920, 1080
which = sys.argv[1] if len(sys.argv) > 1 else "room"
sb, camkw = getattr(scenes, which)()
b = 32 if which == "room" else 8
ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **camkw))
ctx.set_option(A.OPT_MEGAKERNEL_FORM, 1084)
for spp in (8, 32):
    for flags, tag in ((0, "megakernel 8x4"), (A.FLAG_WAVEFRONT, "wavefront")):
        best = 1e30
        for rep in range(4):
            ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=b, flags=flags); ctx.synchronize()
            if rep: best = min(best, ctx.stats()["last_ms"])
        print(f"{which} {spp} spp {tag}: {best:.2f} ms = {w*h*spp/best/1e3:.1f} Mpaths/s", flush=True)
