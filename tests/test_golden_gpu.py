"""GPU parity against the committed golden vectors (tests/golden/reference_vectors.npz, generated from the reference's
own CPU path tracer): closest-hit records bit for bit, accumulation images within the north-star tolerance. Needs no
oracle library at run time — this is the pin that travels."""
import os

import numpy as np
import pytest

import golden_cases as G
from sparkle_b200 import _abi as A
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

pytestmark = pytest.mark.gpu
GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")


@pytest.fixture(scope="module")
def golden():
    return dict(np.load(GOLDEN))


@pytest.mark.parametrize("name", G.SCENES)
def test_hits_and_images_match_reference_vectors(name, golden):
    sb, camkw, rays, bounce = G.small_scene(name)
    desc = sb.finalize()
    ctx = CudaContext(G.IMG_W, G.IMG_H)
    ctx.upload_scene(desc)
    for flags in (0, A.FLAG_EXACT_TRAVERSAL, A.FLAG_THREAD_PER_RAY):
        hits = ctx.trace_closest(rays, flags)
        for f in ("prim", "face"):
            assert np.array_equal(hits[f], golden[f"{name}_hit_{f}"]), (name, f, flags)
        for f in ("t", "u", "v"):
            assert np.array_equal(hits[f].view(np.uint32), golden[f"{name}_hit_{f}"].view(np.uint32)), (name, f, flags)
    ctx.set_camera(orbit_camera(G.IMG_W, G.IMG_H, **camkw))
    ctx.render(1, G.IMG_SPP, max_bounce=bounce)
    gpu = ctx.resolve(A.RESOLVE_LINEAR_F32, total_samples=1)  # the raw sum, like the golden accumulation
    ref = golden[f"{name}_accum"].astype(np.float64)
    rel = np.abs(gpu[..., :3] - ref[..., :3]).sum(-1) / np.maximum(ref[..., :3].sum(-1), 1e-3)
    assert rel.mean() <= 1e-3, (name, rel.mean())
    assert (rel > 1e-4).mean() < 0.02, (name, (rel > 1e-4).mean())
    if name == "testscene":
        ctx.reset_accumulation()
        ctx.render(0, 3, max_bounce=bounce, accum_mode=A.ACCUM_RUNNING_AVERAGE)
        ra = ctx.resolve(A.RESOLVE_LINEAR_F32)
        ref = golden["testscene_running_average"].astype(np.float64)
        rel = np.abs(ra[..., :3] - ref[..., :3]).sum(-1) / np.maximum(ref[..., :3].sum(-1), 1e-3)
        assert rel.mean() <= 1e-3
    ctx.close()
