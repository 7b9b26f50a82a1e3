import sys, time
import os; sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
for name in ("city", "testscene", "room"):
    sb, _ = getattr(scenes, name)()
    d = sb.finalize()
    ctx = CudaContext(64, 64); ctx.set_option(A.OPT_VERBOSE, 1)
    for k in range(3):
        t0 = time.perf_counter(); ctx.upload_scene(d); print(name, "upload ms", round((time.perf_counter() - t0) * 1e3, 2), flush=True)
    ctx.close()
