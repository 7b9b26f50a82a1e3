import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..')); sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'tests'))
import numpy as np
from sparkle_b200 import _abi as A
from sparkle_b200.device import CudaContext
from sparkle_b200.geometry import Xoshiro128Plus
import oracle_lib as O
from test_render_gpu import build
f32 = np.float32
name = sys.argv[1] if len(sys.argv) > 1 else 'testscene'
w, h = 96, 64
sb, desc, cam = build(name, w, h)
orc = O.Oracle('reference')
ctx = CudaContext(w, h); ctx.upload_scene(desc); ctx.set_camera(cam)
out = {}
for dbg in (1, 2, 3):
    ctx.reset_accumulation(); ctx.render(3, 1, max_bounce=8, flags=dbg << 8, output_limit=1e30)
    out[dbg] = ctx.resolve(A.RESOLVE_LINEAR_F32)
# replicate primary rays on the host (RenderPixel + SetupViewRay in float32)
V = lambda a: np.array(a[:], f32)
pos, up, right, mu, mv, ll = V(cam.position), V(cam.up), V(cam.right), V(cam.max_u), V(cam.max_v), V(cam.lower_left)
pw, ph = f32(1) / f32(w - 1), f32(1) / f32(h - 1)
nbad = {1: 0, 2: 0, 3: 0}; shown = 0
for j in range(h):
    for i in range(w):
        rng = Xoshiro128Plus((j * w + i + 3 * w * h) & 0xFFFFFFFF)
        u = f32((f32(i) + rng.unit()) * pw); v = f32((f32(j) + rng.unit()) * ph)
        dx, dy = rng.unit_disk(); dx = f32(dx * f32(cam.aperture_radius)); dy = f32(dy * f32(cam.aperture_radius))
        o = (pos + (dx * right + dy * up)).astype(f32)
        t = ((ll + u * mu).astype(f32) + (v * mv).astype(f32)).astype(f32)
        d = (t - o).astype(f32); nrm = f32(np.sqrt(f32(d[0] * d[0]) + f32(f32(d[1] * d[1]) + f32(d[2] * d[2])))); d = (d / nrm).astype(f32)
        ray = np.zeros(8, f32); ray[0:3] = o; ray[4:7] = d
        o12 = np.zeros(12, f32)
        prim = orc.lib.orc_intersect_full(desc, ray.ctypes.data, o12.ctypes.data_as(A.fp))
        if prim == 0xFFFFFFFF: continue
        refs = {1: o12[4:7], 2: o12[7:10], 3: np.array([o12[10], o12[11], o12[0]], f32)}
        for dbg in (1, 2, 3):
            g = out[dbg][j, i, :3]
            if not np.allclose(g, refs[dbg], rtol=1e-5, atol=1e-6):
                nbad[dbg] += 1
                if shown < 14 and prim in (4, 5):
                    shown += 1; print('pix', (i, j), 'prim', prim, 'dbg', dbg, 'gpu', g, 'ref', refs[dbg])
print('bad normal/tangent/uvt', nbad)
