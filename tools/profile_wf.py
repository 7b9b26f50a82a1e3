"""Short command for ncu: a few wavefront iterations on city (1920x1080, 2 spp, 1M slots)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
name = sys.argv[1] if len(sys.argv) > 1 else "city"
w, h = 1920, 1080
sb, camkw = getattr(scenes, name)()
ctx = CudaContext(w, h)
ctx.upload_scene(sb.finalize())
ctx.set_camera(orbit_camera(w, h, **camkw))
ctx.render(0, int(os.environ.get("SPK_SPP", 2)), max_bounce=8 if name != "room" else 32, flags=A.FLAG_WAVEFRONT)
ctx.synchronize()
print("ms", ctx.stats()["last_ms"])
ctx.close()
