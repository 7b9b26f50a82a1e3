import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
w, h, spp = 3840, 2160, 16
sb, camkw = scenes.city()
ctx = CudaContext(w, h)
ctx.set_stream(torch.cuda.current_stream().cuda_stream)
ctx.upload_scene(sb.finalize())
ctx.set_camera(orbit_camera(w, h, **camkw))
for slots in (1 << 20, 1 << 21, 1 << 22, 1 << 23, 1 << 24):
    ctx.set_option(A.OPT_WAVEFRONT_SLOTS, slots)
    best = 1e9
    for k in range(2):
        ctx.reset_accumulation()
        e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
        e0.record(); ctx.render(0, spp, flags=A.FLAG_WAVEFRONT); e1.record(); torch.cuda.synchronize()
        best = min(best, e0.elapsed_time(e1))
    print("slots", slots, "ms", round(best, 1), "Mpaths/s", round(w * h * spp / best / 1e3, 1), flush=True)
