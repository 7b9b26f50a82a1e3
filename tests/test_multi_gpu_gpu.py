"""Multi-GPU behind the C ABI (include/spk_cuda.h: SpkGroup; csrc/spk_group.cu), on hardware.

Rank g of G renders samples {g, g+G, ...} into its own accumulator; ONE kernel per GPU sums its rows of all accumulators
over peer memory in rank order, tone-maps and gathers at the root. Checked here:
  * the linear frame equals the per-rank accumulators added in rank order on the host, bit for bit (fixed reduce order);
  * it equals a single-context render of the same sample set to 1e-6 (float association only);
  * the tone-mapped fp16 frame equals spk_cuda_resolve of a context holding the same sums, bit for bit;
  * the same through torchrun with one process per GPU (CUDA IPC peers + the flag protocol) when the box has >= 2 GPUs.
A local group works with every context on one GPU too (ordering is stream events, nothing spins), so the first three run
on a 1-GPU box; the IPC form spins on flags and is only ever run with one rank per GPU."""
import os
import subprocess
import sys

import numpy as np
import pytest

from sparkle_b200 import _abi as A
from sparkle_b200 import multi_gpu, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import orbit_camera

pytestmark = pytest.mark.gpu
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))


def _device_count():
    import torch
    return torch.cuda.device_count()


def _render_ranks(world, w, h, spp, devices):
    sb, camkw = scenes.testscene()
    desc = sb.finalize()
    cam = orbit_camera(w, h, **camkw)
    ctxs = []
    for r in range(world):
        c = CudaContext(w, h, device=devices[r])
        c.upload_scene(desc)
        c.set_camera(cam)
        first, count, stride = multi_gpu.sample_slice(spp, r, world)
        c.render(first, count, sample_stride=stride, max_bounce=8)
        ctxs.append(c)
    return ctxs, (sb, desc), cam  # the builder owns the host arrays the descriptor points into


@pytest.mark.parametrize("world", [2, 3])
def test_local_group_equals_host_sum_and_single_context(world):
    w, h, spp = 320, 180, 7
    ndev = _device_count()
    devices = [r % ndev for r in range(world)]
    ctxs, (sb, desc), cam = _render_ranks(world, w, h, spp, devices)
    # per-rank accumulators (sums) as the host sees them
    parts = [c.resolve(A.RESOLVE_LINEAR_F32, total_samples=1) for c in ctxs]
    group = multi_gpu.Group.local(ctxs)
    lin = group.resolve(A.RESOLVE_LINEAR_F32, spp)
    expect = parts[0].copy()
    for p in parts[1:]:
        expect = expect + p  # float32, rank order: the kernel's order
    expect = expect / np.float32(spp)
    assert np.array_equal(lin.view(np.uint32), expect.view(np.uint32))
    # one context, all samples: same set of samples, another association
    one = CudaContext(w, h, device=devices[0])
    one.upload_scene(desc)
    one.set_camera(cam)
    one.render(0, spp, max_bounce=8)
    single = one.resolve(A.RESOLVE_LINEAR_F32)
    assert np.allclose(lin, single, rtol=1e-6, atol=1e-6), float(np.abs(lin - single).max())
    # tone-mapped output: the same per-pixel arithmetic as spk_cuda_resolve (ACES, fp16, row flip)
    f16 = group.resolve(A.RESOLVE_TONEMAP_F16, spp)
    srgb = group.resolve(A.RESOLVE_TONEMAP_SRGB8, spp)
    # a context whose accumulator holds exactly the rank-ordered sums: feed them through the ray-free path
    import torch
    ptr, nbytes = one.accum_ptr()
    one.synchronize()
    summed = parts[0].copy()
    for p in parts[1:]:
        summed = summed + p
    class View:
        __cuda_array_interface__ = {"shape": (nbytes // 4,), "typestr": "<f4", "data": (ptr, False), "version": 3,
                                    "strides": None}
    torch.as_tensor(View(), device=f"cuda:{devices[0]}").copy_(torch.from_numpy(summed.reshape(-1)))
    torch.cuda.synchronize()
    assert np.array_equal(one.resolve(A.RESOLVE_TONEMAP_F16, total_samples=spp).view(np.uint16), f16.view(np.uint16))
    assert np.array_equal(one.resolve(A.RESOLVE_TONEMAP_SRGB8, total_samples=spp), srgb)
    # a second frame through the same group: reset, render, resolve (the events order frame 2 after frame 1's reads)
    for r, c in enumerate(ctxs):
        c.reset_accumulation()
        first, count, stride = multi_gpu.sample_slice(spp, r, world)
        c.render(first, count, sample_stride=stride, max_bounce=8)
    again = group.resolve(A.RESOLVE_LINEAR_F32, spp)
    assert np.array_equal(again, lin)
    group.close()
    for c in ctxs + [one]:
        c.close()


def test_ipc_group_under_torchrun():
    if _device_count() < 2:
        pytest.skip("needs >= 2 GPUs: the IPC form spins on flags, one rank per GPU")
    world = 2
    out = os.path.join(ROOT, "gpurun_out")
    os.makedirs(out, exist_ok=True)
    cmd = [sys.executable, "-m", "torch.distributed.run", "--nnodes=1", f"--nproc-per-node={world}", "--master-addr",
           "127.0.0.1", "--master-port", "29531", os.path.join(ROOT, "tests", "ipc_group_worker.py")]
    r = subprocess.run(cmd, capture_output=True, text=True, timeout=600, cwd=ROOT)
    assert r.returncode == 0, r.stdout[-2000:] + r.stderr[-4000:]
    assert "ipc group ok" in r.stdout
