"""Small run of every kernel family (written for compute-sanitizer, which is closed on this pool; useful as a quick check): wavefront (sorted queue too) and megakernel renders of a small
city, a ray batch through the three trace entry points, a device BVH build, resolve."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import numpy as np
from sparkle_b200 import _abi as A, geometry, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
w, h = 160, 90
sb, camkw = scenes.city(num_meshes=3, tris_per_mesh=500, num_instances=12)
ctx = CudaContext(w, h); ctx.upload_scene(sb.finalize()); ctx.set_camera(orbit_camera(w, h, **dict(camkw, radius=30.0)))
ctx.render(0, 2, max_bounce=8); mega = ctx.resolve(A.RESOLVE_LINEAR_F32)
for sort in (0, 4):
    ctx.set_option(A.OPT_WAVEFRONT_SORT, sort)
    for slots in (1024, 1 << 20):
        ctx.set_option(A.OPT_WAVEFRONT_SLOTS, slots)
        ctx.reset_accumulation(); ctx.render(0, 2, max_bounce=8, flags=A.FLAG_WAVEFRONT)
        wave = ctx.resolve(A.RESOLVE_LINEAR_F32)
        assert np.allclose(wave, mega, rtol=2e-6, atol=1e-6)
rays = geometry.random_rays(50_001, seed=4, extent=20.0)
a = ctx.trace_closest(rays); b = ctx.trace_closest_noalloc(rays); c = ctx.trace_closest(rays, A.FLAG_THREAD_PER_RAY)
assert np.array_equal(a, b) and np.array_equal(a, c)
ctx.resolve(A.RESOLVE_TONEMAP_F16); ctx.close()
print("all kernel families ran; results agree")
