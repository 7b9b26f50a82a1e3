"""A/B of alternative builds of libsparkle_cuda (SPK_CUDA_LIB=<file in sparkle_b200/>): TestScene and room at 1080p, the
measured megakernel form, cost-ordered strips warm; best of 4 device times."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
for name, (w, h, b, spp) in {"testscene": (1920, 1080, 8, 64), "room": (1920, 1080, 32, 32)}.items():
    sb, camkw = getattr(scenes, name)()
    ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **camkw))
    ctx.set_option(A.OPT_MEGAKERNEL_FORM, 1084)
    best = 1e30
    for rep in range(6):
        ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=b); ctx.synchronize()
        if rep > 1: best = min(best, ctx.stats()["last_ms"])
    print(f"{os.environ.get('SPK_CUDA_LIB','default')}: {name} {spp} spp {best:.3f} ms = {w*h*spp/best/1e3:.0f} Mpaths/s", flush=True)
    ctx.close()
