"""Megakernel forms A/B on the bench scenes (device-resident, CUDA events inside the library): for each scene and form,
best-of-3 ms for SPP samples per pixel. usage: r2_forms_ab.py [scene:forms ...]   e.g. testscene:1162,2162,2084
Forms: include/spk_cuda.h SPK_OPT_MEGAKERNEL_FORM. Images of all forms of a scene are compared bit for bit."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

SPP = int(os.environ.get("SPK_SPP", 32))
CFG = {"testscene": (1920, 1080, 8), "room": (1920, 1080, 32), "city": (3840, 2160, 8)}
jobs = sys.argv[1:] or ["testscene:1162,2162,2084", "room:1084,2084,2162"]
for job in jobs:
    name, forms = job.split(":")
    w, h, bounce = CFG[name]
    sb, camkw = getattr(scenes, name)()
    ctx = CudaContext(w, h)
    ctx.upload_scene(sb.finalize())
    ctx.set_camera(orbit_camera(w, h, **camkw))
    first = None
    for form in [int(f) for f in forms.split(",")]:
        ctx.set_option(A.OPT_MEGAKERNEL_FORM, form)
        best = 1e30
        for rep in range(4):
            ctx.reset_accumulation()
            ctx.render(0, SPP, max_bounce=bounce)
            ctx.synchronize()
            if rep:
                best = min(best, ctx.stats()["last_ms"])
        img = ctx.resolve(A.RESOLVE_LINEAR_F32)
        same = "" if first is None else (" same bits" if np.array_equal(img, first) else " IMAGE DIFFERS")
        first = img if first is None else first
        print(f"{name} {w}x{h} {SPP} spp form {form}: {best:.3f} ms = {w * h * SPP / best / 1e3:.0f} Mpaths/s{same}", flush=True)
    ctx.close()
