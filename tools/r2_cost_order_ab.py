import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
for name,(w,h,b) in {"testscene":(1920,1080,8),"room":(1920,1080,32)}.items():
    sb, camkw = getattr(scenes, name)()
    ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **camkw))
    ref=None
    for spp in (32, 256):
      for order in (0, 1):
        ctx.set_option(A.OPT_COST_ORDERED_STRIPS, order)
        best=1e30
        for rep in range(4):
            ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=b); ctx.synchronize()
            if rep: best=min(best, ctx.stats()["last_ms"])
        img=ctx.resolve(A.RESOLVE_LINEAR_F32)
        key=(name,spp)
        same = "" if ref is None or ref[0]!=key else (" same bits" if np.array_equal(img, ref[1]) else " DIFFERS")
        if ref is None or ref[0]!=key: ref=(key,img)
        print(f"{name} {spp} spp cost-order {order}: {best:.3f} ms = {w*h*spp/best/1e3:.0f} Mpaths/s{same}", flush=True)
    ctx.close()
