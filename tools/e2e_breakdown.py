"""Where the end-to-end step goes: host wall time of each C-ABI call (scene upload, camera, render+sync, resolve)."""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
for name, w, h, spp, bounce in (("testscene", 1920, 1080, 32, 8), ("room", 1920, 1080, 16, 32), ("city", 3840, 2160, 4, 8)):
    sb, camkw = getattr(scenes, name)()
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    ctx = CudaContext(w, h)
    for rep in range(3):
        t = [time.perf_counter()]
        ctx.upload_scene(desc); t.append(time.perf_counter())
        ctx.set_camera(cam); ctx.reset_accumulation(); t.append(time.perf_counter())
        ctx.render(0, spp, max_bounce=bounce); ctx.synchronize(); t.append(time.perf_counter())
        ctx.resolve(A.RESOLVE_TONEMAP_F16); t.append(time.perf_counter())
        d = [round((b - a) * 1e3, 2) for a, b in zip(t, t[1:])]
        print(name, "rep", rep, "upload/camera/render/resolve ms", d, flush=True)
    ctx.close()
