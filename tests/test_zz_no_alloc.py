"""spk_cuda_trace_closest_noalloc (no counterpart in the reference: an entry point for callers that must not allocate):
the ray batch call with its staging buffers and ray ticket in the library's static device array. Through the C ABI, its
hits must equal spk_cuda_trace_closest's bit for bit (the tolerance of every traversal test here), on two calls in a row
that share the array: the second is a smaller, different batch, so anything left over from the first would show.
Collected last on purpose."""
import ctypes as C

import numpy as np
import pytest

from sparkle_b200 import _abi as A
from sparkle_b200 import geometry, scenes
from sparkle_b200.device import HIT_DTYPE, CudaContext

pytestmark = pytest.mark.gpu


def bits(x):
    return np.ascontiguousarray(x).view(np.uint32)


def assert_same_hits(got, want, what):
    for f in ("prim", "face"):
        bad = np.nonzero(got[f] != want[f])[0]
        assert bad.size == 0, f"{what}: {bad.size} rays differ in {f}; first {bad[:5]}"
    for f in ("t", "u", "v"):
        bad = np.nonzero(bits(got[f]) != bits(want[f]))[0]
        assert bad.size == 0, f"{what}: {bad.size} rays differ in {f} bits; first {bad[:5]}"


def test_noalloc_matches_trace_closest_twice():
    sb, _ = scenes.city(num_meshes=4, tris_per_mesh=800, num_instances=24)
    desc = sb.finalize()
    ctx = CudaContext(64, 64)
    ctx.upload_scene(desc)

    first = geometry.random_rays(A.NOALLOC_MAX_RAYS, seed=11, extent=20.0)  # the whole array
    first[:, 2] = np.abs(first[:, 2]) * 0.3 + 0.1
    second = geometry.random_rays(50_001, seed=12, extent=20.0)  # ragged, smaller: stale rays / hits lie behind it
    second[:, 2] = np.abs(second[:, 2]) * 0.3 + 0.1

    for k, (rays, flags) in enumerate([(first, 0), (second, 0), (second, A.FLAG_EXACT_TRAVERSAL), (first[:7], 0)]):
        want = ctx.trace_closest(rays, flags)
        hits = np.zeros(len(rays) + 1, HIT_DTYPE)
        hits["prim"][-1] = 0xABCD1234  # guard record behind the output: must stay untouched
        got = ctx.trace_closest_noalloc(rays, flags, hits=hits[:-1])
        assert_same_hits(got, want, f"call {k}")
        assert hits["prim"][-1] == 0xABCD1234
        assert (want["prim"] != A.INVALID_ID).mean() > 0.05
    ctx.close()


def test_noalloc_limits_and_errors():
    sb, _ = scenes.testscene()
    ctx = CudaContext(32, 32)
    lib = ctx.lib
    rays = np.zeros((4, 8), np.float32)
    hits = np.zeros(4, HIT_DTYPE)
    # no scene yet
    assert lib.spk_cuda_trace_closest_noalloc(ctx.handle, rays.ctypes.data, 4, hits.ctypes.data, 0) == A.ERR_NO_SCENE
    ctx.upload_scene(sb.finalize())
    # beyond the static array: refused before anything is touched (the pointers are never read)
    rc = lib.spk_cuda_trace_closest_noalloc(ctx.handle, rays.ctypes.data, A.NOALLOC_MAX_RAYS + 1, hits.ctypes.data, 0)
    assert rc == A.ERR_INVALID_ARGUMENT
    assert b"limit" in lib.spk_cuda_last_error() and str(A.NOALLOC_MAX_RAYS).encode() in lib.spk_cuda_last_error()
    # flags the variant does not implement
    rc = lib.spk_cuda_trace_closest_noalloc(ctx.handle, rays.ctypes.data, 4, hits.ctypes.data, A.FLAG_COUNTERS)
    assert rc == A.ERR_INVALID_ARGUMENT
    assert lib.spk_cuda_trace_closest_noalloc(ctx.handle, None, 4, hits.ctypes.data, 0) == A.ERR_INVALID_ARGUMENT
    assert lib.spk_cuda_trace_closest_noalloc(ctx.handle, None, 0, None, 0) == A.OK
    # and a good call still works afterwards
    rays[:, 0:3], rays[:, 4:7] = (0, 0, 10), (0, 0, -1)
    got = ctx.trace_closest_noalloc(rays)
    assert_same_hits(got, ctx.trace_closest(rays), "after errors")
    ctx.close()
