"""Probe: would two half-pools on two streams overlap wf_shade (DRAM / latency-bound) with wf_extend (issue / L1-bound)?
Two contexts on one GPU render half the samples each from two host threads (the C ABI releases the GIL) against one
context rendering all of them. Wall clock around synchronised calls."""
import os, sys, threading, time
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
which = sys.argv[1] if len(sys.argv) > 1 else "city"
w, h, spp, b = (3840, 2160, 16, 8) if which == "city" else (1920, 1080, 32, 32)
sb, camkw = getattr(scenes, which)()
desc = sb.finalize()
ctxs = []
for k in range(2):
    c = CudaContext(w, h); c.upload_scene(desc); c.set_camera(orbit_camera(w, h, **camkw)); ctxs.append(c)
def run(c, first, count):
    c.reset_accumulation(); c.render(first, count, max_bounce=b, flags=A.FLAG_WAVEFRONT); c.synchronize()
for rep in range(3):
    t0 = time.perf_counter(); run(ctxs[0], 0, spp); t1 = time.perf_counter()
    th = [threading.Thread(target=run, args=(ctxs[k], k * (spp // 2), spp // 2)) for k in range(2)]
    t2 = time.perf_counter()
    for t in th: t.start()
    for t in th: t.join()
    t3 = time.perf_counter()
    print(f"{which}: one context {spp} spp {1e3*(t1-t0):.1f} ms; two contexts x {spp//2} spp concurrently {1e3*(t3-t2):.1f} ms", flush=True)
