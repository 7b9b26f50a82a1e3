"""The two launches whose DRAM traffic bench.py quotes (roofline.traffic): the C2 frame as bench.py launches it (TestScene
1920x1080, ONE 256-spp launch of the measured megakernel form) and the C3 batch (soup-1M, 64 Mi incoherent rays). Run under
  ncu --metrics dram__bytes_read.sum,dram__bytes_write.sum,lts__t_bytes.sum,gpu__time_duration.sum --clock-control none --csv
by tools/r2_dram_traffic.sh; tools/r2_traffic_json.py turns the CSV into profiles/r2_dram_traffic.json."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

w, h = 1920, 1080
sb, camkw = scenes.testscene()
ctx = CudaContext(w, h)
ctx.upload_scene(sb.finalize())
ctx.set_camera(orbit_camera(w, h, **camkw))
# the form is forced: under ncu every launch is replayed and serialised, so spk_cuda_choose_tracer's timings mean nothing
# there. 1084 is what it picks for this scene in a plain run (bench.py prints config.megakernel_form).
FORM = int(os.environ.get("SPK_FORM", 1084))
ctx.set_option(A.OPT_MEGAKERNEL_FORM, FORM)
print("form", (0, FORM))
for k in range(2):
    ctx.reset_accumulation()
    ctx.render(0, 256, max_bounce=8)
    ctx.synchronize()
    print("c2 256 spp ms", ctx.stats()["last_ms"])
ctx.close()
n = 64 << 20
sb, _ = scenes.soup(1_000_000)
ctx = CudaContext(64, 64)
ctx.upload_scene(sb.finalize())
g = torch.Generator(device="cuda").manual_seed(2)
rays = torch.zeros((n, 8), dtype=torch.float32, device="cuda")
rays[:, 0:3] = torch.rand((n, 3), generator=g, device="cuda") * 3.0 - 1.5
d = torch.randn((n, 3), generator=g, device="cuda")
rays[:, 4:7] = d / d.norm(dim=1, keepdim=True)
hits = torch.empty((n, 5), dtype=torch.float32, device="cuda")
for k in range(2):
    ctx.trace_closest_device(rays.data_ptr(), n, hits.data_ptr(), 0)
    ctx.synchronize()
    print("c3 ms", ctx.stats()["last_ms"])
ctx.close()
