"""TEST INFRASTRUCTURE: the fixed, seeded inputs behind tests/golden/reference_vectors.npz and the function that
evaluates them on an oracle. tests/golden/make_golden.py runs evaluate() on the REFERENCE (oracle/_ref) and commits the
result; tests/test_oracle_cpu.py runs it on the port and compares bit for bit; tests/test_golden_gpu.py feeds the same
inputs to the CUDA path through the C ABI."""
import numpy as np

from sparkle_b200 import _abi as A
from sparkle_b200 import geometry, scenes
from sparkle_b200.scene import orbit_camera

RNG_SEEDS = [0, 1, 42, 12345, 921599, 0x7FFFFFFF, 0xFFFFFFFF]
LIBM_OPS = {"sin": 0, "cos": 1, "pow": 2, "sqrt": 3}
IMG_W, IMG_H, IMG_SPP = 48, 32, 2


def small_scene(name):
    """(SceneBuilder, camera kwargs, rays, max_bounce) of the golden scenes: small enough for a fixture, covering spheres,
    instanced meshes, textures (sRGB, unorm, normal maps), dielectrics, emissive stop, closed-room high bounce."""
    if name == "testscene":
        sb, cam = scenes.testscene()
        rays = geometry.random_rays(4096, seed=3, extent=14.0)
        rays[:, 2] = np.abs(rays[:, 2]) * 0.5 + 0.05
        return sb, cam, rays, 8
    if name == "soup":
        sb, cam = scenes.soup(5000, seed=1, size=0.08)
        return sb, cam, geometry.random_rays(4096, seed=2, extent=1.5), 8
    if name == "city":
        sb, cam = scenes.city(num_meshes=3, tris_per_mesh=400, num_instances=12)
        rays = geometry.random_rays(4096, seed=4, extent=15.0)
        rays[:, 2] = np.abs(rays[:, 2]) * 0.3 + 0.1
        return sb, dict(cam, radius=25.0), rays, 8
    if name == "room":
        sb, cam = scenes.room(num_objects=12, tex=16)
        rays = geometry.random_rays(4096, seed=5, extent=4.5)
        rays[:, 2] = np.abs(rays[:, 2]) * 0.5 + 0.3
        return sb, cam, rays, 32
    raise KeyError(name)


SCENES = ["testscene", "soup", "city", "room"]


def kat_vectors(n=64, seed=77):
    g = np.random.default_rng(seed)
    a = g.normal(size=(n, 3)).astype(np.float32) * np.float32(3.0)
    b = g.normal(size=(n, 3)).astype(np.float32)
    m = g.normal(size=(n, 3, 4)).astype(np.float32)
    return a, b, m


def kat_textures():
    g = np.random.default_rng(5)
    rgba8 = g.integers(0, 256, size=(7, 5, 4), dtype=np.uint8)
    f16 = g.uniform(0, 4, size=(6, 6, 4)).astype(np.float16)
    f32 = g.uniform(0, 4, size=(4, 9, 4)).astype(np.float32)
    uv = g.uniform(-1.5, 2.5, size=(48, 2)).astype(np.float32)
    uv[:4] = [[0, 0], [1, 1], [0.5, 0.5], [-0.25, 1.75]]
    return [(rgba8, A.FMT_RGBA8_UNORM), (rgba8, A.FMT_RGBA8_SRGB), (f16, A.FMT_RGBA16F), (f32, A.FMT_RGBA32F)], uv


def kat_sky():
    g = np.random.default_rng(6)
    faces = [g.uniform(0, 3, size=(8, 8, 4)).astype(np.float16) for _ in range(6)]
    dirs = g.normal(size=(64, 3)).astype(np.float32)
    dirs /= np.linalg.norm(dirs, axis=1, keepdims=True).astype(np.float32)
    dirs[:6] = [[1, 0, 0], [-1, 0, 0], [0, 1, 0], [0, -1, 0], [0, 0, 1], [0, 0, -1]]
    return faces, dirs


def sky_desc(color=(0.5, 0.7, 1.0), faces=None, fmt=A.FMT_RGBA16F):
    d = A.SkyDesc()
    d.has_sky = 1
    d.has_map = 0 if faces is None else 1
    d.color[:] = color
    if faces is not None:
        for f in range(6):
            d.faces[f].pixels = faces[f].ctypes.data
            d.faces[f].height, d.faces[f].width = faces[f].shape[:2]
            d.faces[f].format = fmt
    return d


def kat_surface_inputs(n=96, seed=8):
    g = np.random.default_rng(seed)
    normal = g.normal(size=(n, 3)).astype(np.float32)
    normal /= np.linalg.norm(normal, axis=1, keepdims=True).astype(np.float32)
    tangent = np.cross(normal, g.normal(size=(n, 3))).astype(np.float32)
    tangent /= np.linalg.norm(tangent, axis=1, keepdims=True).astype(np.float32)
    d = g.normal(size=(n, 3)).astype(np.float32)
    d /= np.linalg.norm(d, axis=1, keepdims=True).astype(np.float32)
    # incoming direction mostly against the normal (front-face hits), a few from behind
    flip = (np.einsum("ij,ij->i", d, normal) > 0) & (np.arange(n) % 8 != 0)
    d[flip] = -d[flip]
    uv = g.uniform(0, 1, size=(n, 2)).astype(np.float32)
    return d, normal, tangent, uv


def camera_floats(cam):
    """SpkCamera as its 23 floats"""
    return np.frombuffer(bytes(cam), np.float32).copy()


def host_math_inputs(n=256, seed=91):
    """seeded TRS / glTF-node / orbit-camera parameters behind the hostmath_* golden arrays (all float32-exact)"""
    g = np.random.default_rng(seed)
    f = np.float32
    trs_in = np.concatenate([g.uniform(-20, 20, (n, 3)), g.uniform(-3.2, 3.2, (n, 3)), g.uniform(0.1, 5, (n, 3))], 1).astype(f)
    trs_in[::4, 7:9] = trs_in[::4, 6:7]   # uniform scales
    trs_in[::7, 3] = 0                     # axis-aligned rotations
    trs_in[0, 3:6] = 0
    trs_in[1] = [0, 0, -0.001, 0, 0, 0, 50, 50, 0.001]  # TestScene floor (UsdRoundTripTest.cpp:40-45)
    node_in = np.concatenate([g.uniform(-20, 20, (n, 3)), g.normal(size=(n, 4))], 1).astype(f)
    node_in[::3, 3:] /= np.linalg.norm(node_in[::3, 3:], axis=1, keepdims=True)
    node_in[0, 3:] = [0, 0, 0, 1]
    cam_in = np.stack([g.uniform(-5, 5, n), g.uniform(-5, 5, n), g.uniform(-5, 5, n), g.uniform(1, 60, n),
                       g.uniform(-89, 89, n), g.uniform(-180, 180, n), g.uniform(0.015, 0.08, n), np.full(n, 0.024),
                       g.uniform(1, 22, n), np.ones(n), g.integers(16, 4000, n), g.integers(16, 2200, n)], 1).astype(f)
    cam_in[0] = [0, 0, 0, 25, 10, -20, 0.035, 0.024, 22, 1, 1920, 1080]  # TestScene camera (UsdRoundTripTest.cpp:28-33)
    cam_in[1] = [0, 0, 0, 25, 10, -20, 0.035, 0.024, 22, 1, 1280, 720]
    return trs_in, node_in, cam_in


def evaluate_host_math(fn_trs, fn_node, fn_camera):
    """fn_trs(t, e, s) -> (fwd, inv); fn_node(t, q) -> (fwd, inv); fn_camera(w, h, **kw) -> SpkCamera. Evaluated on the
    reference (orc.kat_*) for the golden file and on the product's host library in tests/test_host_cpu.py."""
    trs_in, node_in, cam_in = host_math_inputs()
    out = {}
    r = [fn_trs(x[0:3], x[3:6], x[6:9]) for x in trs_in]
    out["hostmath_trs_fwd"] = np.stack([a for a, _ in r]).astype(np.float32)
    out["hostmath_trs_inv"] = np.stack([b for _, b in r]).astype(np.float32)
    r = [fn_node(x[0:3], x[3:7]) for x in node_in]
    out["hostmath_node_fwd"] = np.stack([a for a, _ in r]).astype(np.float32)
    out["hostmath_node_inv"] = np.stack([b for _, b in r]).astype(np.float32)
    out["hostmath_camera"] = np.stack([camera_floats(fn_camera(
        int(x[10]), int(x[11]), center=tuple(float(v) for v in x[0:3]), radius=float(x[3]), pitch=float(x[4]),
        yaw=float(x[5]), focal_length=float(x[6]), sensor_height=float(x[7]), aperture=float(x[8]),
        exposure=float(x[9]))) for x in cam_in])
    return out


def sky_scene():
    """TestScene under a seeded RGBA16F cube-map sky (a14 with has_map, SkyRenderProxy.cpp:50-57, Image.cpp:298-357)"""
    sb, cam = scenes.testscene()
    g = np.random.default_rng(14)
    faces = []
    for f in range(6):
        y, x = np.mgrid[0:32, 0:32] / 31.0
        base = np.stack([0.3 + 0.7 * x, 0.2 + 0.8 * y, 0.5 + 0.5 * np.sin(6.0 * (x + y) + f), np.ones_like(x)], -1)
        faces.append((base * (0.6 + 0.25 * f) + g.uniform(0, 0.3, base.shape)).astype(np.float16))
    sb.set_sky(color=(0.5, 0.7, 1.0), faces=faces, fmt=A.FMT_RGBA16F)
    return sb, cam


WRAP_W, WRAP_H, WRAP_FRAME = 3840, 2160, 517  # C4's frame: p + 517*W*H crosses 2^32 at pixel 6,762,496


def wrap_probe_pixels():
    """pixels on either side of the 32-bit seed wrap of frame 517 at 3840x2160 (and the two neighbours of the crossing)"""
    cross = (1 << 32) - WRAP_FRAME * WRAP_W * WRAP_H
    ps = [0, 12345, cross - 2, cross - 1, cross, cross + 1, cross + 4321, WRAP_W * WRAP_H - 1]
    return [(p % WRAP_W, p // WRAP_W) for p in ps]


def evaluate(orc):
    """all golden quantities from one oracle -> dict of numpy arrays (floats kept as float32: compare the bits)"""
    out = {}
    out["rng_state"] = np.stack([orc.rng_state(s) for s in RNG_SEEDS])
    out["rng_units"] = np.stack([orc.rng_units(s, 16) for s in RNG_SEEDS])
    out["unit_disk"] = np.stack([orc.unit_disk(s) for s in RNG_SEEDS])
    a, b, m = kat_vectors()
    for op, name in enumerate(["dot", "squared_norm", "normalized", "cross", "transform_direction", "transform_point"]):
        res = np.stack([orc.kat_vec(op, a[k], b[k], m[k]) for k in range(len(a))])
        out[f"vec_{name}"] = res[:, :1] if op < 2 else res
    texs, uv = kat_textures()
    out["texture"] = np.stack([np.stack([orc.texture(px, fmt, u, v) for (u, v) in uv]) for px, fmt in texs])
    faces, dirs = kat_sky()
    sd_map, sd_grad = sky_desc(faces=faces), sky_desc()
    out["sky_map"] = np.stack([orc.sky(sd_map, d) for d in dirs])
    out["sky_gradient"] = np.stack([orc.sky(sd_grad, d) for d in dirs])
    for name in SCENES:
        sb, camkw, rays, bounce = small_scene(name)
        desc = sb.finalize()
        hits, _ = orc.trace_closest(desc, rays, threads=1)
        for f in ("t", "u", "v", "prim", "face"):
            out[f"{name}_hit_{f}"] = np.ascontiguousarray(hits[f])
        full = [orc.intersect_full(desc, r) for r in rays[:256]]
        out[f"{name}_full_prim"] = np.array([p for p, _ in full], np.uint32)
        out[f"{name}_full"] = np.stack([np.where(p != A.INVALID_ID, o, 0).astype(np.float32) for p, o in full])
        cam = orbit_camera(IMG_W, IMG_H, **camkw)
        acc, _ = orc.render(desc, cam, IMG_W, IMG_H, 1, IMG_SPP, max_bounce=bounce, threads=4)
        out[f"{name}_accum"] = acc
        if name in ("testscene", "room"):
            d, n, t, uvs = kat_surface_inputs()
            wi, thr, draws = [], [], []
            for k in range(len(d)):
                w, th, dr = orc.sample_surface(desc, k % len(sb.materials), 1000 + k, d[k], n[k], t[k], uvs[k])
                wi.append(w), thr.append(th), draws.append(dr)
            out[f"{name}_surface_wi"] = np.stack(wi)
            out[f"{name}_surface_throughput"] = np.stack(thr)
            out[f"{name}_surface_draws"] = np.array(draws, np.uint32)
        if name == "testscene":
            ra, cum = orc.render(desc, cam, IMG_W, IMG_H, 0, 3, max_bounce=bounce, accum_mode=A.ACCUM_RUNNING_AVERAGE,
                                 threads=4)
            out["testscene_running_average"] = ra
            out["testscene_tonemap_f16"] = orc.resolve(acc, A.RESOLVE_TONEMAP_F16, IMG_SPP, exposure=1.25).view(np.uint16)
            # H6: single pixel samples of C4's frame 517 at 3840x2160, seeds on both sides of the 2^32 wrap
            cam4k = orbit_camera(WRAP_W, WRAP_H, **camkw)
            out["testscene_wrap_samples"] = np.stack([orc.sample_pixel(desc, cam4k, WRAP_W, WRAP_H, i, j, WRAP_FRAME)[:4]
                                                      for (i, j) in wrap_probe_pixels()])
    # a14 with a cube map: image of TestScene under a seeded RGBA16F cube sky
    sb, camkw = sky_scene()
    desc = sb.finalize()
    acc, _ = orc.render(desc, orbit_camera(IMG_W, IMG_H, **camkw), IMG_W, IMG_H, 1, IMG_SPP, max_bounce=8, threads=4)
    out["skyscene_accum"] = acc
    return out
