import sys, os
R = os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."); sys.path.insert(0, R); sys.path.insert(0, os.path.join(R, "tests"))
import numpy as np
import oracle_lib as O, golden_cases as G
from sparkle_b200 import _abi as A, app, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
ref = O.Oracle("reference"); port = O.Oracle("port")
def cube():
    w, h, spp = 192, 108, 4
    sb, camkw = G.sky_scene(); desc = sb.finalize(); cam = orbit_camera(w, h, **camkw)
    ctx = CudaContext(w, h); ctx.upload_scene(desc); ctx.set_camera(cam); ctx.render(0, spp, max_bounce=8)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32); ctx.close()
    r, _ = ref.render(desc, cam, w, h, 0, spp, max_bounce=8); p, _ = port.render(desc, cam, w, h, 0, spp, max_bounce=8)
    r /= spp; p /= spp
    f = lambda a, b: float(np.abs(a[..., :3] - b[..., :3]).mean())
    print("gpu-ref", f(gpu, r), "gpu-port", f(gpu, p), "ref-port", f(r, p), flush=True)
cube()
which = sys.argv[1] if len(sys.argv) > 1 else "all"
w, h = 96, 64
if which in ("all", "gpus"):
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp 4 --spp 2 --spatial_denoise false --gpus 2 --gpu_devices 0,0"); a.run(100); a.read_linear(); a.close()
    print("after gpus"); cube()
if which in ("all", "usd"):
    from sparkle_b200 import usd_export
    sb, camkw = scenes.testscene(); path = usd_export.write_usda(sb, "/tmp/ts.usda", camera=camkw)
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp 4 --spp 2 --spatial_denoise false --scene {path}"); a.run(); img = a.read_linear()
    r2, _ = ref.render(a.scene, a.camera, w, h, 0, 4, accum_mode=A.ACCUM_RUNNING_AVERAGE); a.close()
    print("after usd"); cube()
if which in ("all", "gltf"):
    from sparkle_b200 import gltf_export
    sb, cam = scenes.city(num_meshes=4, tris_per_mesh=800, num_instances=24); path = gltf_export.write_gltf(sb, "/tmp/city.gltf")
    a = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp 4 --spp 1 --spatial_denoise false --scene {path} --gltf_instancing true"); a.run(); a.read_linear()
    r2, _ = ref.render(a.scene, a.camera, w, h, 0, 4, accum_mode=A.ACCUM_RUNNING_AVERAGE); a.close()
    print("after gltf"); cube()
