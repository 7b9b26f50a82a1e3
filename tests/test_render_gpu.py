"""GPU parity of the path tracer: spk_cuda_render / spk_cuda_resolve vs the oracle's CPURenderer on the same scene,
camera and sample seeds. Tolerance (BASELINE.json north_star): mean relative per-pixel error <= 1e-3 at equal spp with
identical sample streams; PSNR reported. Against the reference build the images are expected to be the same (the device
computes sin, cos and pow the way glibc does, csrc/spk_glibc_math.cuh): asserted to 1e-7 / 99.9 % identical pixels."""
import numpy as np
import pytest

import parity_report as PR
from sparkle_b200 import _abi as A
from sparkle_b200 import scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

pytestmark = pytest.mark.gpu

MEAN_REL_TOL = 1e-3


def image_metrics(gpu, ref, name=None):
    """aggregate metrics + worst 8x8 tile; with a name also the signed-difference image and a report line
    (tests/parity_report.py, SURVEY.md section 4: never aggregate-only)"""
    if name:
        return PR.report(name, gpu, ref)
    return PR.metrics(gpu, ref)[0]


def build(name, w, h):
    if name == "testscene":
        sb, cam = scenes.testscene()
    elif name == "city":
        sb, cam = scenes.city(num_meshes=4, tris_per_mesh=800, num_instances=24)
        cam = dict(cam, radius=30.0)
    elif name == "room":
        sb, cam = scenes.room(num_objects=20, tex=32)
    desc = sb.finalize()
    return sb, desc, orbit_camera(w, h, **cam)


@pytest.mark.parametrize("name,bounce", [("testscene", 8), ("city", 8), ("room", 32)])
def test_one_sample_per_pixel(name, bounce, oracle):
    w, h = 96, 64
    sb, desc, cam = build(name, w, h)
    ref, _ = oracle.render(desc, cam, w, h, first_sample=3, sample_count=1, max_bounce=bounce)
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(first_sample=3, sample_count=1, max_bounce=bounce)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32)
    m = image_metrics(gpu, ref)
    print(name, m)
    # single samples: a path that takes a different branch shows up as a wholly different pixel
    assert m["frac_gt_1e4"] < 0.01, m
    assert np.array_equal(gpu[..., 3] > 0, ref[..., 3] > 0) or (np.mean((gpu[..., 3] > 0) != (ref[..., 3] > 0)) < 0.01)
    st = ctx.stats()
    assert st["paths"] == w * h and st["rays"] >= w * h
    ctx.close()


@pytest.mark.parametrize("name,bounce,spp", [("testscene", 8, 32), ("city", 8, 16)])
def test_image_tolerance(name, bounce, spp, oracle):
    w, h = 128, 72
    sb, desc, cam = build(name, w, h)
    ref, _ = oracle.render(desc, cam, w, h, first_sample=0, sample_count=spp, max_bounce=bounce)
    ref = ref / spp
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(first_sample=0, sample_count=spp, max_bounce=bounce)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32)
    m = image_metrics(gpu, ref, f"image_tolerance_{name}_{spp}spp")
    assert m["mean_rel"] <= MEAN_REL_TOL, m
    assert m["psnr"] > 40.0, m
    if oracle.kind == "reference":  # same libm on both sides since csrc/spk_glibc_math.cuh: the same image
        assert m["mean_rel"] <= 1e-7 and m["frac_identical"] > 0.999, m
    # exact (never-culled) traversal gives the same image as the culled one
    ctx.reset_accumulation()
    ctx.render(first_sample=0, sample_count=spp, max_bounce=bounce, flags=A.FLAG_EXACT_TRAVERSAL)
    exact = ctx.resolve(A.RESOLVE_LINEAR_F32)
    assert np.array_equal(exact, gpu)
    ctx.close()


def test_sample_split_and_stride():
    """sample s of pixel p depends only on (p, s): launches and GPU-style strided splits reproduce the same sum"""
    w, h = 64, 48
    sb, desc, cam = build("testscene", w, h)
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(0, 8)
    whole = ctx.resolve(A.RESOLVE_LINEAR_F32)
    ctx.reset_accumulation()
    ctx.render(0, 3)
    ctx.render(3, 5)
    assert ctx.accum_sample_count() == 8
    assert np.array_equal(ctx.resolve(A.RESOLVE_LINEAR_F32), whole)  # same order of additions: bit-exact
    parts = []
    for g in range(2):
        ctx.reset_accumulation()
        ctx.render(g, 4, sample_stride=2)
        parts.append(ctx.resolve(A.RESOLVE_LINEAR_F32, total_samples=1))
    both = (parts[0] + parts[1]) / 8
    assert np.allclose(both, whole, rtol=1e-5, atol=1e-6)
    ctx.close()


def test_running_average_and_resolve_formats(oracle):
    w, h = 64, 48
    sb, desc, cam = build("testscene", w, h)
    acc, cum = oracle.render(desc, cam, w, h, 0, 6, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(0, 4, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    ctx.render(4, 2, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32)
    m = image_metrics(gpu, acc)
    assert cum == 6 and m["mean_rel"] <= MEAN_REL_TOL, m
    with pytest.raises(A.SpkError):
        ctx.render(6, 1, accum_mode=A.ACCUM_SUM)  # mode change without reset
    # tone-mapped fp16, rows flipped (CPURenderer.cpp:565-576): oracle applied to the GPU's own accumulator
    f16 = ctx.resolve(A.RESOLVE_TONEMAP_F16)
    ref16 = oracle.resolve(gpu, A.RESOLVE_TONEMAP_F16, 1, exposure=1.0, accum_mode=A.ACCUM_RUNNING_AVERAGE)
    assert np.array_equal(f16.view(np.uint16), ref16.view(np.uint16))
    srgb = ctx.resolve(A.RESOLVE_TONEMAP_SRGB8)
    assert srgb.shape == (h, w, 4) and srgb[..., 3].min() == 255 and srgb[..., :3].max() > 100
    ctx.close()


@pytest.mark.parametrize("name,bounce,spp", [("testscene", 8, 8), ("city", 8, 4), ("room", 32, 4)])
def test_wavefront_equals_megakernel(name, bounce, spp, oracle):
    """SPK_FLAG_WAVEFRONT: the same samples through the queue tracer; only the order of the per-pixel sum differs"""
    w, h = 128, 72
    sb, desc, cam = build(name, w, h)
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(first_sample=2, sample_count=spp, max_bounce=bounce)
    mega = ctx.resolve(A.RESOLVE_LINEAR_F32)
    st_m = ctx.stats()
    for slots in (1024, 1048576):  # pool smaller than one sample of the image / larger than the whole job
        ctx.set_option(A.OPT_WAVEFRONT_SLOTS, slots)
        ctx.reset_accumulation()
        ctx.reset_stats()
        ctx.render(first_sample=2, sample_count=spp, max_bounce=bounce, flags=A.FLAG_WAVEFRONT)
        wave = ctx.resolve(A.RESOLVE_LINEAR_F32)
        st_w = ctx.stats()
        assert np.allclose(wave, mega, rtol=2e-6, atol=1e-6), np.abs(wave - mega).max()
        assert st_w["paths"] == w * h * spp
        assert st_w["rays"] == st_m["rays"]  # the same paths, segment for segment
    # SPK_OPT_WAVEFRONT_SORT: the ray queue sorted by origin cell + octant before it is traced; the same samples again
    ctx.set_option(A.OPT_WAVEFRONT_SORT, 4)
    ctx.reset_accumulation()
    ctx.reset_stats()
    ctx.render(first_sample=2, sample_count=spp, max_bounce=bounce, flags=A.FLAG_WAVEFRONT)
    sorted_queue = ctx.resolve(A.RESOLVE_LINEAR_F32)
    assert np.allclose(sorted_queue, mega, rtol=2e-6, atol=1e-6), np.abs(sorted_queue - mega).max()
    assert ctx.stats()["rays"] == st_m["rays"]
    ctx.set_option(A.OPT_WAVEFRONT_SORT, 0)
    ref, _ = oracle.render(desc, cam, w, h, first_sample=2, sample_count=spp, max_bounce=bounce)
    m = image_metrics(wave, ref / spp)
    if name == "room":
        # depth 32 in a closed room: a path that takes another branch near a threshold is a different pixel, and dark
        # pixels make that a large relative error; bound how many such pixels there are instead
        assert m["frac_gt_1e4"] < 0.01 and m["psnr"] > 40.0, m
    else:
        assert m["mean_rel"] <= MEAN_REL_TOL, m
    with pytest.raises(A.SpkError):
        ctx.render(0, 1, accum_mode=A.ACCUM_RUNNING_AVERAGE, flags=A.FLAG_WAVEFRONT)
    ctx.close()


def test_spatial_denoise(oracle):
    """a15 (SPK_FLAG_SPATIAL_DENOISE): g-buffer sample, 8-neighbour fill of invalid pixels, running average"""
    import oracle_lib as O
    # the port: the reference's scatter is a race between row workers once its pool has more than one (the pool is
    # created once per process); the port is the one-worker order, pinned against the reference in test_oracle_cpu.py
    orc = O.Oracle("port")
    w, h = 96, 64
    sb, desc, cam = build("room", w, h)
    ref, cum = orc.render(desc, cam, w, h, 0, 3, max_bounce=32, accum_mode=A.ACCUM_RUNNING_AVERAGE, threads=1,
                          flags=A.FLAG_SPATIAL_DENOISE)
    plain, _ = orc.render(desc, cam, w, h, 0, 3, max_bounce=32, accum_mode=A.ACCUM_RUNNING_AVERAGE, threads=1)
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(cam)
    ctx.render(0, 2, max_bounce=32, accum_mode=A.ACCUM_RUNNING_AVERAGE, flags=A.FLAG_SPATIAL_DENOISE)
    ctx.render(2, 1, max_bounce=32, accum_mode=A.ACCUM_RUNNING_AVERAGE, flags=A.FLAG_SPATIAL_DENOISE)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32)
    assert ctx.accum_sample_count() == 3 and cum == 3
    changed = (ref != plain).any(-1)
    assert changed.any(), "the scene must exercise the fill"
    m = image_metrics(gpu, ref)
    assert m["frac_gt_1e4"] < 0.01, m
    # the filled pixels carry the neighbour's colour exactly as in the reference
    rel = np.abs(gpu[..., :3] - ref[..., :3]).sum(-1) / np.maximum(ref[..., :3].sum(-1), 1e-3)
    assert (rel[changed] < 1e-4).mean() > 0.97, rel[changed]
    assert np.array_equal(gpu[..., 3][changed], ref[..., 3][changed])
    ctx.close()


def test_cost_ordered_strips_do_not_change_the_image():
    """the sorted megakernel launches its strips of tiles in the order of their cost in the previous launch (longest
    first): a schedule only - every launch order gives the same image bit for bit, for both CTA shapes, also when the
    order was learnt on another camera and another scene of the same resolution"""
    w, h, spp = 640, 360, 8
    sb, camkw = scenes.testscene()
    desc = sb.finalize()
    ctx = CudaContext(w, h)
    ctx.upload_scene(desc)
    ctx.set_camera(orbit_camera(w, h, **camkw))
    for form in (1084, 1162):
        ctx.set_option(A.OPT_MEGAKERNEL_FORM, form)
        ctx.set_option(A.OPT_COST_ORDERED_STRIPS, 0)
        ctx.reset_accumulation()
        ctx.render(0, spp, max_bounce=8)
        plain = ctx.resolve(A.RESOLVE_LINEAR_F32)
        ctx.set_option(A.OPT_COST_ORDERED_STRIPS, 1)
        for launch in range(3):  # the first learns the costs, the next two run in cost order
            ctx.reset_accumulation()
            ctx.render(0, spp, max_bounce=8)
            assert np.array_equal(ctx.resolve(A.RESOLVE_LINEAR_F32), plain), (form, launch)
    # a stale order (learnt above) on another view and another scene: still only a schedule
    ctx.set_camera(orbit_camera(w, h, **dict(camkw, yaw=70.0, pitch=35.0)))
    ctx.reset_accumulation()
    ctx.render(0, spp, max_bounce=8)
    moved = ctx.resolve(A.RESOLVE_LINEAR_F32)
    ctx.set_option(A.OPT_COST_ORDERED_STRIPS, 0)
    ctx.reset_accumulation()
    ctx.render(0, spp, max_bounce=8)
    assert np.array_equal(ctx.resolve(A.RESOLVE_LINEAR_F32), moved)
    ctx.close()
