"""GPU parity: closest-hit (prim, face, t-bits, u, v) of the CUDA traversal vs the oracle, through the C ABI
(spk_cuda_trace_closest with HOST buffers). Bit-exact: integer ids equal, float fields compared as uint32."""
import numpy as np
import pytest

from sparkle_b200 import _abi as A
from sparkle_b200 import geometry, scenes
from sparkle_b200.device import CudaContext

pytestmark = pytest.mark.gpu


def bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def assert_hits_equal(gpu, ref, what):
    for f in ("prim", "face"):
        bad = np.nonzero(gpu[f] != ref[f])[0]
        assert bad.size == 0, f"{what}: {bad.size} rays differ in {f}; first {bad[:5]} gpu={gpu[bad[:5]]} ref={ref[bad[:5]]}"
    for f in ("t", "u", "v"):
        bad = np.nonzero(bits(gpu[f]) != bits(ref[f]))[0]
        assert bad.size == 0, f"{what}: {bad.size} rays differ in {f} bits; first {bad[:5]} gpu={gpu[bad[:5]]} ref={ref[bad[:5]]}"


def scene_rays(name, n):
    if name == "testscene":
        sb, _ = scenes.testscene()
        rays = geometry.random_rays(n, seed=3, extent=14.0)
        rays[:, 2] = np.abs(rays[:, 2]) * 0.5 + 0.05
    elif name == "soup":
        sb, _ = scenes.soup(20000, seed=1, size=0.06)
        rays = geometry.random_rays(n, seed=2, extent=1.5)
    elif name == "city":
        sb, _ = scenes.city(num_meshes=4, tris_per_mesh=800, num_instances=24)
        rays = geometry.random_rays(n, seed=4, extent=20.0)
        rays[:, 2] = np.abs(rays[:, 2]) * 0.3 + 0.1
    elif name == "room":
        sb, _ = scenes.room(num_objects=20, tex=32)
        rays = geometry.random_rays(n, seed=5, extent=4.5)
        rays[:, 2] = np.abs(rays[:, 2]) * 0.5 + 0.3
    return sb, rays


@pytest.mark.parametrize("name", ["testscene", "soup", "city", "room"])
def test_closest_hit_bit_exact(name, oracle):
    sb, rays = scene_rays(name, 200_000)
    desc = sb.finalize()
    ref, _ = oracle.trace_closest(desc, rays)
    ctx = CudaContext(64, 64)
    ctx.upload_scene(desc)
    exact = ctx.trace_closest(rays, A.FLAG_EXACT_TRAVERSAL)
    assert_hits_equal(exact, ref, f"{name} exact traversal")
    culled = ctx.trace_closest(rays)
    assert_hits_equal(culled, ref, f"{name} culled traversal")
    counted = ctx.trace_closest(rays, A.FLAG_COUNTERS)
    assert_hits_equal(counted, ref, f"{name} counted traversal")
    st = ctx.stats()
    assert st["node_visits"] > 0 and st["kernel_launches"] == 3
    assert (ref["prim"] != A.INVALID_ID).mean() > 0.05
    ctx.close()


def test_degenerate_rays(oracle):
    """axis-parallel directions (d.x == 0 drops every mesh hit), rays starting on surfaces, misses"""
    sb, _ = scenes.testscene()
    desc = sb.finalize()
    rays = np.zeros((6, 8), np.float32)
    rays[0, 0:3], rays[0, 4:7] = (0, 0, 10), (0, 0, -1)
    rays[1, 0:3], rays[1, 4:7] = (5, 4, 10), (0, 0, -1)       # d.x == 0 over a mesh
    rays[2, 0:3], rays[2, 4:7] = (-30, 0, 2), (1, 0, 0)
    rays[3, 0:3], rays[3, 4:7] = (0, 0, 50), (0, 0, 1)        # miss
    rays[4, 0:3], rays[4, 4:7] = (3, 3, 0), (0.6, 0, 0.8)     # starts on the floor
    rays[5, 0:3], rays[5, 4:7] = (0, -30, 2), (0, 1, 0)
    ref, _ = oracle.trace_closest(desc, rays, threads=1)
    ctx = CudaContext(64, 64)
    ctx.upload_scene(desc)
    assert_hits_equal(ctx.trace_closest(rays, A.FLAG_EXACT_TRAVERSAL), ref, "degenerate exact")
    assert_hits_equal(ctx.trace_closest(rays), ref, "degenerate culled")
    ctx.close()


def test_errors():
    ctx = CudaContext(64, 64)
    rays = geometry.random_rays(16)
    with pytest.raises(A.SpkError) as e:
        ctx.trace_closest(rays)
    assert e.value.code == A.ERR_NO_SCENE
    sb, _ = scenes.soup(100)
    d = sb.finalize()
    d.abi_version = 99
    with pytest.raises(A.SpkError) as e:
        ctx.upload_scene(d)
    assert e.value.code == A.ERR_INVALID_ARGUMENT
    ctx.close()


def test_update_instances_validation_and_extent(oracle):
    """spk_cuda_update_instances: a stale material index is rejected (it would index the device material table out of
    bounds), and instances moved far outside the uploaded world bounds still give culled == exact == oracle: the
    t-culling slack is re-derived from the rebuilt TLAS root."""
    sb, _ = scenes.city(num_meshes=3, tris_per_mesh=400, num_instances=12)
    desc = sb.finalize()
    ctx = CudaContext(64, 64)
    ctx.upload_scene(desc)
    bad = sb.primitive_array()
    bad[2].material = len(sb.materials) + 5
    with pytest.raises(A.SpkError) as e:
        ctx.update_instances(bad, len(sb.prims), sb.tlas.desc())
    assert e.value.code == A.ERR_INVALID_ARGUMENT and "material" in str(e.value)
    # move every instance 400 units out (the upload's extent was ~30): rebuild the TLAS host-side, update, trace
    from sparkle_b200.geometry import trs
    for p in sb.prims[1:]:
        t, eul, scl = p["trs"]
        t2 = (t[0] + 400.0, t[1] - 250.0, t[2])
        fwd, inv = trs(t2, eul, scl)
        lo, hi = sb.meshes[p["mesh"]].bounds()
        corners = np.array([[x, y, z] for x in (lo[0], hi[0]) for y in (lo[1], hi[1]) for z in (lo[2], hi[2])], np.float32)
        wc = corners @ fwd[:, :3].T + fwd[:, 3]
        p.update(fwd=fwd, inv=inv, lo=wc.min(axis=0), hi=wc.max(axis=0), trs=(t2, eul, scl))
    moved = sb.finalize()  # same meshes, new transforms and TLAS
    ctx.update_instances(sb.primitive_array(), len(sb.prims), sb.tlas.desc())
    rays = geometry.random_rays(50_000, seed=9, extent=20.0)
    rays[:, 0] += 400.0
    rays[:, 1] -= 250.0
    rays[:, 2] = np.abs(rays[:, 2]) * 0.3 + 0.1
    ref, _ = oracle.trace_closest(moved, rays)
    assert (ref["prim"] != A.INVALID_ID).mean() > 0.05
    assert_hits_equal(ctx.trace_closest(rays, A.FLAG_EXACT_TRAVERSAL), ref, "moved, exact")
    assert_hits_equal(ctx.trace_closest(rays), ref, "moved, culled")
    ctx.close()


def test_pinned_host_buffers_take_the_pipelined_path():
    """spk_cuda_trace_closest with page-locked host buffers overlaps copy-in, traversal and copy-out in pieces of 2 Mi rays:
    the hits are those of the one-shot path (pageable buffers), bit for bit, including a ragged last piece"""
    import ctypes as C
    sb, _ = scenes.testscene()
    ctx = CudaContext(64, 64)
    ctx.upload_scene(sb.finalize())
    n = (5 << 20) + 12345
    rays = geometry.random_rays(n, seed=21, extent=14.0)
    plain = ctx.trace_closest(rays)
    lib = ctx.lib
    p_rays, p_hits = C.c_void_p(), C.c_void_p()
    A.check(lib.spk_cuda_host_alloc(C.byref(p_rays), rays.nbytes))
    A.check(lib.spk_cuda_host_alloc(C.byref(p_hits), n * 20))
    C.memmove(p_rays.value, rays.ctypes.data, rays.nbytes)
    A.check(lib.spk_cuda_trace_closest(ctx.handle, p_rays, n, p_hits, 0))
    piped = np.frombuffer((C.c_uint8 * (n * 20)).from_address(p_hits.value), dtype=plain.dtype).copy()
    for f in ("prim", "face"):
        assert np.array_equal(piped[f], plain[f])
    for f in ("t", "u", "v"):
        assert np.array_equal(piped[f].view(np.uint32), plain[f].view(np.uint32))
    assert ctx.stats()["last_ms"] > 0
    lib.spk_cuda_host_free(p_rays)
    lib.spk_cuda_host_free(p_hits)
    ctx.close()
