"""Megakernel vs wavefront (SPK_FLAG_WAVEFRONT) on the bench scenes: CUDA-event time of the whole render call.
usage: wf_compare.py [scene ...]   scenes: testscene city room ; env SPK_SPP, SPK_WF_SLOTS_LIST"""
import os, sys, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
import torch
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

names = sys.argv[1:] or ["testscene"]
spp = int(os.environ.get("SPK_SPP", 32))
for name in names:
    if name == "testscene":
        sb, camkw = scenes.testscene(); w, h, bounce = 1920, 1080, 8
    elif name == "city":
        sb, camkw = scenes.city(); w, h, bounce = 1920, 1080, 8
    else:
        sb, camkw = scenes.room(); w, h, bounce = 1920, 1080, 32
    t0 = time.time()
    desc = sb.finalize()
    print(name, "scene build s", round(time.time() - t0, 1), flush=True)
    ctx = CudaContext(w, h)
    ctx.set_stream(torch.cuda.current_stream().cuda_stream)
    ctx.upload_scene(desc)
    ctx.set_camera(orbit_camera(w, h, **camkw))
    def run(flags):
        best = 1e9
        for k in range(3):
            ctx.reset_accumulation(); ctx.reset_stats()
            e0, e1 = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
            e0.record(); ctx.render(0, spp, max_bounce=bounce, flags=flags); e1.record()
            torch.cuda.synchronize()
            best = min(best, e0.elapsed_time(e1))
        st = ctx.stats()
        return best, st
    ms, st = run(0)
    img_m = ctx.resolve(A.RESOLVE_LINEAR_F32)
    print(f"{name} megakernel {spp} spp: {ms:.2f} ms  {w*h*spp/ms/1e3:.0f} Mpaths/s  {st['rays']/ms/1e3:.0f} Mrays/s rays/path {st['rays']/st['paths']:.2f}", flush=True)
    for slots in os.environ.get("SPK_WF_SLOTS_LIST", "524288,1048576,2097152,4194304").split(","):
        ctx.set_option(A.OPT_WAVEFRONT_SLOTS, int(slots))
        ms, st = run(A.FLAG_WAVEFRONT)
        img_w = ctx.resolve(A.RESOLVE_LINEAR_F32)
        err = float(np.abs(img_w - img_m).max())
        print(f"{name} wavefront slots {slots}: {ms:.2f} ms  {w*h*spp/ms/1e3:.0f} Mpaths/s  {st['rays']/ms/1e3:.0f} Mrays/s launches {st['kernel_launches']} max|diff| {err:.2e}", flush=True)
    ctx.close()
