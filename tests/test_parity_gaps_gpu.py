"""GPU parity gates that round 1 did not have ("parity gaps"), all through the C ABI, all against the
reference's own CPU path tracer (oracle/_ref) where it is built, else the pinned port:

  C5   closed room, depth 32 + Russian roulette, textured: vs the CPU path tracer at equal spp. The GPU must equal the port
       run in "device libm" mode (glibc's sin / cos / pow, which the device restates, and Schlick's pow(x, 5) as the device's
       exact product) to accumulation rounding, and the reference itself to 1e-7 (it read 4.3e-3 while the device evaluated
       sin / cos / pow in double: the gate separates such a libm share from "same operations in the same order")
  H6   32-bit seed wrap: frame 517 at 3840x2160 (C4's configuration) crosses 2^32 inside the frame
  a14  cube-map sky on the device: image of a scene under a cube sky, and a miss-ray probe (empty scene: every pixel is
       SkyRenderProxy::Evaluate of its camera ray)
  a16  sRGB8 resolve: bit-exact against the port, within one code of the reference
  a15  running average past max_spp: the capped moving average of CameraRenderProxy::Update (CameraRenderProxy.cpp:70-73)
Every image gate writes a signed-difference image and the worst 8x8 tile (tests/parity_report.py)."""
import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
import parity_report as PR
from sparkle_b200 import _abi as A
from sparkle_b200 import scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import SceneBuilder, orbit_camera

pytestmark = pytest.mark.gpu
MEAN_REL_TOL = 1e-3  # BASELINE.json north_star: mean relative per-pixel error at equal spp, identical sample streams


@pytest.fixture()
def port_device_libm():
    if not O.available("port"):
        pytest.skip("oracle/liboracle.so not built")
    port = O.Oracle("port")
    port.set_libm(True)
    yield port
    port.set_libm(False)


def gpu_render(desc, cam, w, h, first, count, bounce, **kw):
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(first_sample=first, sample_count=count, max_bounce=bounce, **kw)
    img = ctx.resolve(A.RESOLVE_LINEAR_F32)
    st = ctx.stats()
    ctx.close()
    return img, st


def test_c5_room_depth32_gate(oracle, port_device_libm):
    """C5 at the north-star metric. 480x270 x 128 spp of the full room (50 objects, 256^2 textures), depth 32."""
    w, h, spp, bounce = 480, 270, 128, 32
    sb, camkw = scenes.room()
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    gpu, st = gpu_render(desc, cam, w, h, 0, spp, bounce)
    ref, _ = oracle.render(desc, cam, w, h, 0, spp, max_bounce=bounce, threads=15)
    same_libm, _ = port_device_libm.render(desc, cam, w, h, 0, spp, max_bounce=bounce, threads=15)
    m_ops = PR.report("c5_room_vs_port_device_libm", gpu, same_libm / spp, dict(spp=spp, bounce=bounce))
    m_ref = PR.report(f"c5_room_vs_{oracle.kind}", gpu, ref / spp,
                      dict(spp=spp, bounce=bounce, rays_per_path=st["rays"] / st["paths"]))
    m_libm = PR.report("c5_room_port_device_libm_vs_reference_cpu_only", same_libm / spp, ref / spp, dict(spp=spp))
    # (1) same operations, same order, same libm: nothing but the odd double-rounding tie may differ. This is the gate
    # that says the kernels compute what the reference computes, at the north-star metric, three decades inside it.
    assert m_ops["mean_rel"] <= 1e-5 and m_ops["frac_identical"] > 0.99, m_ops
    # (2) against the reference CPU path tracer itself. A third of this image is BLACK in the reference (no path of
    # 128 found the ceiling panel; there is no next-event estimation), and a path that an ulp of sin/cos/pow sends to the
    # light instead is then an unbounded relative error: measured CPU-only (port with the device's libm vs the reference,
    # no GPU involved) six such pixels alone give 4.3e-3. So the plain metric is asserted to be nothing but that libm
    # share, and the 1e-3 bar is asserted where a relative error is defined (lit pixels) and as relative L1.
    assert m_ref["mean_rel"] <= 1.25 * m_libm["mean_rel"] + 1e-4, (m_ref, "libm share", m_libm)
    assert m_ref["mean_rel_lit"] <= MEAN_REL_TOL and m_ref["rel_l1"] <= MEAN_REL_TOL, m_ref
    assert m_ref["frac_gt_1e4"] < 0.005 and m_ref["psnr"] > 60.0, m_ref
    # (3) since the device restates glibc's sinf / cosf / powf (csrc/spk_glibc_math.cuh) there is no libm share left to
    # separate out: against the reference build the depth-32 room is the same image (Schlick's pow(x, 5) aside)
    if oracle.kind == "reference":
        assert m_ref["mean_rel"] <= 1e-7 and m_ref["frac_identical"] > 0.999, m_ref


@pytest.mark.parametrize("name", ["testscene", "city"])
def test_gpu_equals_port_with_device_libm(name, port_device_libm):
    """the same separation on the other two scene families: with the libm difference removed the images are identical"""
    w, h, spp = 160, 90, 16
    if name == "testscene":
        sb, camkw = scenes.testscene()
    else:
        sb, camkw = scenes.city(num_meshes=4, tris_per_mesh=800, num_instances=24)
        camkw = dict(camkw, radius=30.0)
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    gpu, _ = gpu_render(desc, cam, w, h, 0, spp, 8)
    same_libm, _ = port_device_libm.render(desc, cam, w, h, 0, spp, max_bounce=8)
    m = PR.report(f"{name}_vs_port_device_libm", gpu, same_libm / spp, dict(spp=spp))
    assert m["mean_rel"] <= 1e-6 and m["frac_identical"] > 0.995, m


def test_h6_seed_wrap_at_4k(oracle, golden_wrap):
    """seed = j*W + i + frame*W*H in 32-bit unsigned arithmetic (CPURenderer.cpp:442-443): at 3840x2160 frame 517 wraps
    2^32 at pixel 6,762,496. Frames 517..519 against the CPU path tracer, the two sides of the wrap separately."""
    w, h, frame = G.WRAP_W, G.WRAP_H, G.WRAP_FRAME
    sb, camkw = scenes.testscene()
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(first_sample=frame, sample_count=1)
    one = ctx.resolve(A.RESOLVE_LINEAR_F32)
    # the committed single-pixel samples of the reference on both sides of the crossing
    for (i, j), want in zip(G.wrap_probe_pixels(), golden_wrap):
        assert np.allclose(one[j, i], want, rtol=1e-5, atol=1e-6), ((i, j), one[j, i], want)
    ctx.render(first_sample=frame + 1, sample_count=2)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32)
    ctx.close()
    ref, _ = oracle.render(desc, cam, w, h, frame, 3, threads=15)
    ref = ref / 3
    cross = ((1 << 32) - frame * w * h) // w  # first row with wrapped seeds
    for part, rows in (("before_wrap", slice(0, cross)), ("after_wrap", slice(cross + 1, h))):
        m = PR.report(f"h6_wrap_4k_{part}", gpu[rows], ref[rows], dict(frames=[frame, frame + 1, frame + 2]))
        assert m["mean_rel"] <= MEAN_REL_TOL and m["frac_gt_1e4"] < 0.01, (part, m)


@pytest.fixture()
def golden_wrap():
    import os
    path = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")
    return np.load(path)["testscene_wrap_samples"]


def test_cube_map_sky_image(oracle):
    """a14 with has_map: TestScene under a seeded RGBA16F cube sky (golden_cases.sky_scene)"""
    w, h, spp = 192, 108, 16
    sb, camkw = G.sky_scene()
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    gpu, _ = gpu_render(desc, cam, w, h, 0, spp, 8)
    ref, _ = oracle.render(desc, cam, w, h, 0, spp, max_bounce=8)
    m = PR.report("a14_cube_sky_testscene", gpu, ref / spp, dict(spp=spp))
    if m["mean_rel"] > MEAN_REL_TOL:  # say which side moved: the device or the reference build
        import oracle_lib as O
        port, _ = O.Oracle("port").render(desc, cam, w, h, 0, spp, max_bounce=8)
        again, _ = gpu_render(desc, cam, w, h, 0, spp, 8)
        m["diag"] = dict(gpu_vs_port=float(np.abs(gpu - port / spp).mean()), ref_vs_port=float(np.abs(ref - port).mean() / spp),
                         gpu_again_vs_gpu=float(np.abs(again - gpu).mean()))
    assert m["mean_rel"] <= MEAN_REL_TOL and m["psnr"] > 40.0, m
    # the wavefront tracer reads the same six faces (entries of the texture table): same samples, other summation order
    wave, st_w = gpu_render(desc, cam, w, h, 0, spp, 8, flags=A.FLAG_WAVEFRONT)
    assert np.allclose(wave, gpu, rtol=2e-6, atol=1e-6), np.abs(wave - gpu).max()
    # the committed golden image of the same scene (the pin that travels)
    import os
    golden = np.load(os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz"))
    small, _ = gpu_render(desc, orbit_camera(G.IMG_W, G.IMG_H, **camkw), G.IMG_W, G.IMG_H, 1, G.IMG_SPP, 8)
    want = golden["skyscene_accum"] / G.IMG_SPP
    rel = np.abs(small[..., :3] - want[..., :3]).sum(-1) / np.maximum(want[..., :3].sum(-1), 1e-3)
    assert rel.mean() <= MEAN_REL_TOL, rel.mean()


@pytest.mark.parametrize("fmt,dtype", [(A.FMT_RGBA16F, np.float16), (A.FMT_RGBA32F, np.float32)])
def test_cube_map_sky_miss_probe(oracle, fmt, dtype):
    """an EMPTY scene under a cube sky: every camera ray misses, so every pixel of a 1-spp image is
    SkyRenderProxy::Evaluate(ray) = Image2DCube::Sample(direction) (SkyRenderProxy.cpp:50-57, Image.cpp:298-357): face
    choice, per-face uv and the face-local bilinear filter, pixel by pixel. A wide lens covers five faces."""
    w, h = 256, 256
    g = np.random.default_rng(3)
    faces = [(g.uniform(0, 4, (16, 16, 4)) * (1 + f)).astype(dtype) for f in range(6)]
    worst = 0.0
    for pitch, yaw in ((10.0, -20.0), (-80.0, 45.0), (85.0, 170.0)):
        sb = SceneBuilder()
        sb.set_sky(color=(0.5, 0.7, 1.0), faces=faces, fmt=fmt)
        desc = sb.finalize()
        cam = orbit_camera(w, h, radius=5.0, pitch=pitch, yaw=yaw, focal_length=0.006)  # ~127 degrees vertical
        gpu, st = gpu_render(desc, cam, w, h, 7, 1, 8)
        ref, _ = oracle.render(desc, cam, w, h, 7, 1, max_bounce=8)
        assert st["rays"] == w * h  # one segment per path: all miss
        assert gpu[..., :3].max() > 1.0 and len(np.unique(np.round(gpu[..., 0], 3))) > 100
        # the lens sample's sin/cos may differ by one ulp: the direction moves by ~1e-7, the filtered colour with it
        err = np.abs(gpu[..., :3] - ref[..., :3]) / np.maximum(ref[..., :3], 1e-3)
        # a direction within an ulp of a face edge or a texel boundary may pick the other side on one of the two: count
        assert (err.max(-1) > 1e-4).mean() < 2e-4, (pitch, yaw, float((err.max(-1) > 1e-4).mean()))
        worst = max(worst, float(np.median(err)))
    assert worst < 1e-6


def test_srgb8_resolve_bits(oracle, port_device_libm):
    """a16, the sRGB8 branch: ACES, then LinearToSrgb = pow(c, 0.45) (Utilities.h:411-414), truncated to 8 bits"""
    w, h = 96, 64
    sb, camkw = scenes.testscene()
    desc = sb.finalize()
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(orbit_camera(w, h, exposure=1.3, **camkw))
    ctx.render(0, 8)
    lin = ctx.resolve(A.RESOLVE_LINEAR_F32, total_samples=1)  # the raw sum
    srgb = ctx.resolve(A.RESOLVE_TONEMAP_SRGB8)
    ctx.close()
    same = port_device_libm.resolve(lin, A.RESOLVE_TONEMAP_SRGB8, 8, exposure=1.3)
    assert np.array_equal(srgb, same), int((srgb != same).sum())
    port_device_libm.set_libm(False)
    ref = port_device_libm.resolve(lin, A.RESOLVE_TONEMAP_SRGB8, 8, exposure=1.3)  # glibc powf, what the reference calls
    diff = np.abs(srgb.astype(np.int32) - ref.astype(np.int32))
    assert diff.max() <= 1 and (diff > 0).mean() < 0.01, (int(diff.max()), float((diff > 0).mean()))


def test_running_average_past_max_spp(oracle):
    """a15 / a17: once the camera's cumulated count reaches max_spp it stays there (CameraRenderProxy.cpp:70-73), so
    DenoisePass keeps lerping with cap/(cap+1): an exponential moving average over fresh seeds, not a running mean.
    Through the C ABI (cumulated_cap) and through the host app (RenderConfig::max_spp), against the CPU path tracer."""
    from sparkle_b200 import app
    w, h, cap, frames = 64, 48, 3, 7
    sb, camkw = scenes.testscene()
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    ref, cum = oracle.render(desc, cam, w, h, 0, frames, accum_mode=A.ACCUM_RUNNING_AVERAGE, cumulated_cap=cap)
    mean, _ = oracle.render(desc, cam, w, h, 0, frames, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert np.abs(ref - mean).max() > 1e-3, "the cap must change the image"
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(0, 2, accum_mode=A.ACCUM_RUNNING_AVERAGE, cumulated_cap=cap)
    ctx.render(2, frames - 2, accum_mode=A.ACCUM_RUNNING_AVERAGE, cumulated_cap=cap)  # crosses the cap inside a call
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32)
    ctx.close()
    m = PR.report("a15_running_average_past_cap", gpu, ref)
    assert m["mean_rel"] <= MEAN_REL_TOL, m
    # the host app renders one sample per frame and keeps rendering after it is ready for the screenshot
    happ = app.HostApp(f"--pipeline cuda --width {w} --height {h} --max_spp {cap} --spp 1 --spatial_denoise false "
                       "--scene testscene")
    for _ in range(frames):
        happ.render_frame()
    host_img = happ.read_linear()
    hdesc, hcam = happ.scene, happ.camera
    href, _ = oracle.render(hdesc, hcam, w, h, 0, frames, accum_mode=A.ACCUM_RUNNING_AVERAGE, cumulated_cap=cap)
    mh = PR.report("a15_host_app_past_cap", host_img, href)
    happ.close()
    assert mh["mean_rel"] <= MEAN_REL_TOL, mh
