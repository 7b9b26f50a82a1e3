"""The phased megakernel is what a launch falls back to when the sorted form cannot pack its path slot (more than 255
bounces or 2^24 samples in one call). Same arithmetic in the same order: the images must be bit-identical."""
import numpy as np
import pytest

from sparkle_b200 import _abi as A, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

pytestmark = pytest.mark.gpu


def test_phased_fallback_equals_sorted_form():
    w, h, spp = 320, 180, 4
    sb, camkw = scenes.testscene()
    ctx = CudaContext(w, h)
    ctx.upload_scene(sb.finalize())
    ctx.set_camera(orbit_camera(w, h, **camkw))
    ctx.render(0, spp, max_bounce=255)  # sorted form
    sorted_form = ctx.resolve(A.RESOLVE_LINEAR_F32)
    launches = ctx.stats()["kernel_launches"]
    ctx.reset_accumulation()
    # no path of this open scene survives 255 bounces of Russian roulette, so the cap does not change any sample
    ctx.render(0, spp, max_bounce=300)  # phased kernel
    assert ctx.stats()["kernel_launches"] > launches
    assert np.array_equal(ctx.resolve(A.RESOLVE_LINEAR_F32), sorted_form)
    ctx.close()
