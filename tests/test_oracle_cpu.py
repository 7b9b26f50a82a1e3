"""CPU tests of the oracle (no GPU): the C++ port (oracle/liboracle.so) against
  - tests/golden/reference_vectors.npz, generated from the reference's OWN CPU path tracer (oracle/_ref, the reference
    translation units compiled in place) by tests/golden/make_golden.py  -> bit for bit;
  - the reference library itself where it is built (this container)         -> bit for bit, larger cases;
  - known answers probed from the vendored XoshiroCpp.hpp;
  - brute force (every primitive, every triangle), which is BVH-topology independent.
"""
import os

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
from sparkle_b200 import _abi as A
from sparkle_b200 import geometry, scenes
from sparkle_b200.scene import orbit_camera

GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")


def bits(x):
    x = np.ascontiguousarray(x)
    return x.view(np.uint32) if x.dtype == np.float32 else x


def assert_same_bits(got, want, what):
    assert got.shape == want.shape and got.dtype == want.dtype, f"{what}: {got.shape}/{got.dtype} vs {want.shape}/{want.dtype}"
    bad = np.nonzero(bits(got).reshape(-1) != bits(want).reshape(-1))[0]
    assert bad.size == 0, f"{what}: {bad.size} of {got.size} elements differ; first at {bad[:5]}: " \
                          f"{got.reshape(-1)[bad[:5]]} vs {want.reshape(-1)[bad[:5]]}"


@pytest.fixture(scope="module")
def port():
    if not O.available("port"):
        pytest.fail("oracle/liboracle.so missing: make -C oracle")
    o = O.Oracle("port")
    assert o.kind == "port"
    return o


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(GOLDEN))


def test_port_matches_reference_golden_vectors(port, golden):
    got = G.evaluate(port)
    # hostmath_*: the reference's scene-graph math, which pins the PRODUCT's host library (tests/test_host_cpu.py)
    assert sorted(got) == sorted(k for k in golden if not k.startswith("hostmath_"))
    for k in sorted(got):
        assert_same_bits(got[k], golden[k], f"golden[{k}]")


@pytest.mark.skipif(not O.available("reference"), reason="oracle/_ref not built (needs /root/reference)")
def test_reference_reproduces_golden_vectors(golden):
    ref = O.Oracle("reference")
    assert ref.kind == "reference"
    got = G.evaluate(ref)
    got.update(G.evaluate_host_math(ref.kat_transform, ref.kat_gltf_node, ref.kat_camera))
    assert sorted(got) == sorted(golden)
    for k in sorted(golden):
        assert_same_bits(got[k], golden[k], f"golden[{k}]")


def test_rng_known_answers(port):
    # SURVEY.md 8a row a1, probed from thirdparty/xoshiro_cpp/XoshiroCpp.hpp
    assert [f"{x:08x}" for x in port.rng_state(0)] == ["7b1dcdaf", "a1b965f4", "8009454f", "724c81ec"]
    u = port.rng_units(12345, 2)
    assert float(u[0]).hex() == "0x1.84151a0000000p-6" and float(u[1]).hex() == "0x1.a33ba40000000p-2"
    # the python generator that seeds the procedural scenes is the same stream
    for seed in G.RNG_SEEDS:
        g = geometry.Xoshiro128Plus(seed)
        mine = np.array([g.unit() for _ in range(16)], np.float32)
        assert_same_bits(mine, port.rng_units(seed, 16), f"python xoshiro seed {seed}")


def test_eigen_association(port):
    """H4: the association the reference's Eigen expressions compile to; a case where the other order differs"""
    a, b = np.array([1e8, 1.0, -1e8], np.float32), np.array([1.0, 1.0, 1.0], np.float32)
    assert port.kat_vec(0, a, b)[0] == np.float32(np.float32(np.float32(1e8) + np.float32(1.0)) + np.float32(-1e8))
    a = np.array([1.0, 3e-4, 3e-4], np.float32)
    want = np.float32(a[0] * a[0]) + np.float32(np.float32(a[1] * a[1]) + np.float32(a[2] * a[2]))
    assert port.kat_vec(1, a)[0] == np.float32(want)


@pytest.mark.parametrize("name", G.SCENES)
def test_bvh_traversal_equals_brute_force(port, name):
    """closest hit is topology independent except for exact-t ties (strict <): BVH, culled BVH and brute force agree"""
    sb, _, rays, _ = G.small_scene(name)
    desc = sb.finalize()
    bvh, c0 = port.trace_closest(desc, rays)
    brute, c1 = port.trace_closest(desc, rays, mode=O.TRACE_BRUTE_FORCE)
    culled, c2 = port.trace_closest(desc, rays, mode=O.TRACE_BVH_CULLED)
    assert_same_bits(bvh["t"], brute["t"], f"{name} t bvh vs brute")
    assert_same_bits(bvh["t"], culled["t"], f"{name} t bvh vs culled")
    tie_free = bvh["prim"] == brute["prim"]
    assert tie_free.mean() > 0.999
    assert np.array_equal(bvh["prim"], culled["prim"]) and np.array_equal(bvh["face"], culled["face"])
    assert c2.node_visits <= c0.node_visits and c0.tri_tests <= c1.tri_tests
    assert (bvh["prim"] != A.INVALID_ID).mean() > 0.05


def test_empty_and_degenerate(port):
    sb = scenes.SceneBuilder()
    sb.set_sky((0.5, 0.7, 1.0))
    desc = sb.finalize()
    rays = geometry.random_rays(8, seed=1)
    hits, _ = port.trace_closest(desc, rays, threads=1)
    assert (hits["prim"] == A.INVALID_ID).all() and (hits["t"] == np.finfo(np.float32).max).all()
    # zero rays
    hits, _ = port.trace_closest(desc, np.zeros((0, 8), np.float32))
    assert len(hits) == 0
    # d.x == 0 drops every mesh hit (world_t = (p.x - o.x) / d.x is NaN/inf); spheres are unaffected
    sb, _ = scenes.testscene()
    desc = sb.finalize()
    r = np.zeros((2, 8), np.float32)
    r[0, 0:3], r[0, 4:7] = (5, 4, 10), (0, 0, -1)   # over the box stand-in: mesh + floor, both dropped
    r[1, 0:3], r[1, 4:7] = (0, 0, 10), (0, 0, -1)   # over the glass sphere
    hits, _ = port.trace_closest(desc, r, threads=1)
    assert hits["prim"][0] == A.INVALID_ID and hits["prim"][1] == 1 and hits["t"][1] == np.float32(6.0)


def test_sample_split_invariance(port):
    """sample s of pixel p depends only on (p, s): any split of the sample range gives the same set of samples"""
    sb, camkw, _, bounce = G.small_scene("testscene")
    desc = sb.finalize()
    w, h = 40, 24
    cam = orbit_camera(w, h, **camkw)
    whole, _ = port.render(desc, cam, w, h, 0, 4, max_bounce=bounce)
    a, _ = port.render(desc, cam, w, h, 0, 2, max_bounce=bounce, sample_stride=2)
    b, _ = port.render(desc, cam, w, h, 1, 2, max_bounce=bounce, sample_stride=2)
    assert np.allclose(a + b, whole, rtol=1e-6, atol=1e-6)
    one, _ = port.render(desc, cam, w, h, 2, 1, max_bounce=bounce)
    px = port.sample_pixel(desc, cam, w, h, 7, 5, 2, max_bounce=bounce)
    assert_same_bits(px[:4], one[5, 7], "sample_pixel vs render")


@pytest.mark.skipif(not O.available("reference"), reason="oracle/_ref not built (needs /root/reference)")
@pytest.mark.parametrize("name,n_rays,spp", [("testscene", 60000, 4), ("room", 30000, 2), ("city", 30000, 2)])
def test_port_equals_reference_larger(port, name, n_rays, spp):
    ref = O.Oracle("reference")
    if name == "testscene":
        sb, camkw = scenes.testscene()
        rays = geometry.random_rays(n_rays, seed=13, extent=14.0)
        bounce = 8
    elif name == "room":
        sb, camkw = scenes.room(num_objects=20, tex=32)
        rays = geometry.random_rays(n_rays, seed=15, extent=4.5)
        bounce = 32
    else:
        sb, camkw = scenes.city(num_meshes=4, tris_per_mesh=800, num_instances=24)
        camkw = dict(camkw, radius=30.0)
        rays = geometry.random_rays(n_rays, seed=14, extent=20.0)
        bounce = 8
    desc = sb.finalize()
    hr, _ = ref.trace_closest(desc, rays)
    hp, _ = port.trace_closest(desc, rays)
    for f in ("prim", "face", "t", "u", "v"):
        assert_same_bits(hp[f], hr[f], f"{name} hit.{f}")
    w, h = 96, 64
    cam = orbit_camera(w, h, **camkw)
    ar, _ = ref.render(desc, cam, w, h, 5, spp, max_bounce=bounce)
    ap, _ = port.render(desc, cam, w, h, 5, spp, max_bounce=bounce)
    assert_same_bits(ap, ar, f"{name} accumulation image")


def test_spatial_denoise_port_equals_reference(tmp_path):
    """a15: SpatialDenoisePixel + temporal lerp (CPURenderer.cpp:452-552); the room scene has invalid pixels to fill.
    The reference scatters from a ParallelFor over rows, so with several workers two rows can race for the same invalid
    pixel; its worker pool is created once per process, so the reference side runs in a fresh process with ONE worker,
    which is the deterministic order the port (and the CUDA gather) reproduce."""
    if not (O.available("port") and O.available("reference")):
        pytest.skip("needs both oracle libraries")
    import subprocess
    import sys
    w, h = 96, 64
    sb, camkw = scenes.room(num_objects=20, tex=32)
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    orc = O.Oracle("port")
    plain, _ = orc.render(desc, cam, w, h, 0, 3, accum_mode=A.ACCUM_RUNNING_AVERAGE, threads=1)
    den, cum = orc.render(desc, cam, w, h, 0, 3, accum_mode=A.ACCUM_RUNNING_AVERAGE, threads=1, flags=A.FLAG_SPATIAL_DENOISE)
    assert cum == 3 and (plain != den).any(), "the scene must exercise the fill"
    out = tmp_path / "ref_denoised.npy"
    code = (
        "import sys, numpy as np\n"
        f"sys.path.insert(0, {os.path.dirname(os.path.abspath(__file__))!r}); sys.path.insert(0, {os.path.dirname(os.path.dirname(os.path.abspath(__file__)))!r})\n"
        "import oracle_lib as O\n"
        "from sparkle_b200 import _abi as A, scenes\n"
        "from sparkle_b200.scene import orbit_camera\n"
        "sb, camkw = scenes.room(num_objects=20, tex=32)\n"
        "img, _ = O.Oracle('reference').render(sb.finalize(), orbit_camera(96, 64, **camkw), 96, 64, 0, 3, "
        "accum_mode=A.ACCUM_RUNNING_AVERAGE, threads=1, flags=A.FLAG_SPATIAL_DENOISE)\n"
        f"np.save({str(out)!r}, img)\n")
    subprocess.check_call([sys.executable, "-c", code])
    assert np.array_equal(den, np.load(out))
    with pytest.raises(RuntimeError):
        O.Oracle("port").render(desc, cam, w, h, 0, 1, flags=A.FLAG_SPATIAL_DENOISE)  # sum mode: not the reference's pass


@pytest.mark.skipif(not O.available("reference"), reason="oracle/_ref not built (needs /root/reference)")
def test_running_average_cap_port_equals_reference(port):
    """cumulated_cap (the camera's max_spp cap, CameraRenderProxy.cpp:70-73): port == reference bit for bit, and the
    capped average differs from the running mean"""
    ref = O.Oracle("reference")
    sb, camkw = scenes.testscene()
    desc = sb.finalize()
    cam = orbit_camera(40, 30, **camkw)
    a, _ = ref.render(desc, cam, 40, 30, 0, 6, accum_mode=A.ACCUM_RUNNING_AVERAGE, cumulated_cap=2, threads=2)
    b, _ = port.render(desc, cam, 40, 30, 0, 6, accum_mode=A.ACCUM_RUNNING_AVERAGE, cumulated_cap=2, threads=2)
    c, _ = port.render(desc, cam, 40, 30, 0, 6, accum_mode=A.ACCUM_RUNNING_AVERAGE, threads=2)
    assert_same_bits(b, a, "capped running average")
    assert np.abs(b - c).max() > 1e-3


def test_device_libm_mode_of_the_port(port):
    """orc_set_libm(1): what the CUDA path computes: glibc's sin / cos / pow like the reference, except Schlick's
    (1 - cos)^5 as an exact product rounded once (one ulp from std::pow(x, 5.f) for ~0.1 % of the arguments); mode 0
    restores the reference's bits."""
    x = np.random.default_rng(5).uniform(0, 6.3, 20000).astype(np.float32)
    base = np.array([port.kat_libm(0, v) for v in x[:2000]])
    sb, camkw = scenes.testscene()
    desc = sb.finalize()
    cam = orbit_camera(48, 32, **camkw)
    plain, _ = port.render(desc, cam, 48, 32, 0, 4, threads=2)
    try:
        port.set_libm(True)
        dev, _ = port.render(desc, cam, 48, 32, 0, 4, threads=2)
    finally:
        port.set_libm(False)
    again, _ = port.render(desc, cam, 48, 32, 0, 4, threads=2)
    assert_same_bits(again, plain, "mode 0 restored")
    rel = np.abs(dev[..., :3] - plain[..., :3]).sum(-1) / np.maximum(plain[..., :3].sum(-1), 1e-3)
    assert rel.mean() < 1e-5, rel.mean()
    assert np.array_equal(base, np.array([port.kat_libm(0, v) for v in x[:2000]]))
