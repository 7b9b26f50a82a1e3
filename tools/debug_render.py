"""debug helper (GPU box): where do 1-spp pixels differ from the oracle? histogram by primary-hit primitive"""
import sys, os
sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..')); sys.path.insert(0, os.path.join(os.path.dirname(__file__), '..', 'tests'))
import numpy as np
from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera
import oracle_lib as O
name = sys.argv[1] if len(sys.argv) > 1 else 'testscene'
bounce = int(sys.argv[2]) if len(sys.argv) > 2 else 8
w, h = 96, 64
from test_render_gpu import build
sb, desc, cam = build(name, w, h)
orc = O.Oracle('reference')
ctx = CudaContext(w, h); ctx.upload_scene(desc); ctx.set_camera(cam)
for b in ([1, 2, bounce] if bounce > 2 else [bounce]):
    ref, _ = orc.render(desc, cam, w, h, 3, 1, max_bounce=b)
    ctx.reset_accumulation(); ctx.render(3, 1, max_bounce=b)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32)
    rel = np.abs(gpu[..., :3] - ref[..., :3]).sum(-1) / np.maximum(ref[..., :3].sum(-1), 1e-3)
    bad = rel > 1e-4
    print('bounce', b, 'bad frac', bad.mean(), 'valid mismatch', ((gpu[..., 3] > 0) != (ref[..., 3] > 0)).mean())
# primary hit per pixel centre
pos = np.array(cam.position[:]); ll = np.array(cam.lower_left[:]); mu = np.array(cam.max_u[:]); mv = np.array(cam.max_v[:])
jj, ii = np.mgrid[0:h, 0:w]
tgt = ll + ((ii + 0.5) / (w - 1))[..., None] * mu + ((jj + 0.5) / (h - 1))[..., None] * mv
d = tgt - pos; d /= np.linalg.norm(d, axis=-1, keepdims=True)
rays = np.zeros((h * w, 8), np.float32); rays[:, 0:3] = pos; rays[:, 4:7] = d.reshape(-1, 3)
hits, _ = orc.trace_closest(desc, rays)
prim = hits['prim'].reshape(h, w).astype(np.int64); prim[prim == 0xFFFFFFFF] = -1
for p in np.unique(prim):
    m = prim == p
    mat = sb.prims[p]['material'] if p >= 0 else -1
    print('prim', p, 'mat', mat, sb.materials[mat]['type'] if p >= 0 else '', 'pixels', m.sum(), 'bad', bad[m].sum())
ys, xs = np.nonzero(bad)
for y, x in list(zip(ys, xs))[:8]:
    print((x, y), 'prim', prim[y, x], 'gpu', gpu[y, x], 'ref', ref[y, x])
