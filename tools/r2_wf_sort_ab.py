"""A/B: wavefront ray sorting (SPK_OPT_WAVEFRONT_SORT, bits per axis of the origin cell in the sort key) on city-10M at 4K
and on the scenes the wavefront is not chosen for. Prints ms per step, Mpaths/s, and the image difference against the
unsorted run (same samples, only the order of the per-pixel atomic sums differs)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

cases = {"city": (3840, 2160, 8, 16), "room": (1920, 1080, 32, 16), "testscene": (1920, 1080, 8, 16)}
which = sys.argv[1].split(",") if len(sys.argv) > 1 else list(cases)
bits_list = [int(b) for b in sys.argv[2].split(",")] if len(sys.argv) > 2 else [0, 3, 4, 5, 6, 7]
for name in which:
    w, h, b, spp = cases[name]
    sb, camkw = getattr(scenes, name)()
    ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **camkw))
    ref = None
    for bits in bits_list:
        ctx.set_option(A.OPT_WAVEFRONT_SORT, bits)
        best = 1e30
        for rep in range(3):
            ctx.reset_accumulation(); ctx.render(0, spp, max_bounce=b, flags=A.FLAG_WAVEFRONT); ctx.synchronize()
            if rep: best = min(best, ctx.stats()["last_ms"])
        img = ctx.resolve(A.RESOLVE_LINEAR_F32)
        if ref is None: ref = img
        d = float(np.abs(img - ref).max())
        st = ctx.stats()
        print(f"{name} {w}x{h} {spp} spp sort bits {bits}: {best:.2f} ms = {w*h*spp/best/1e3:.1f} Mpaths/s; max |diff| vs unsorted {d:.2e}", flush=True)
    ctx.close()
