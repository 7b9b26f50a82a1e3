"""N>1 host logic on CPU: world_size-2 gloo. Each rank renders its interleaved sample subset (with the oracle port
standing in for the device — this is a test of the partitioning and the reduce, not of the kernels), the accumulators are
reduced to rank 0 and compared with the single-rank sum over the same set of sample indices."""
import os
import socket

import numpy as np
import pytest
import torch
import torch.distributed as dist
import torch.multiprocessing as mp

import golden_cases as G
import oracle_lib as O
from sparkle_b200 import multi_gpu
from sparkle_b200.scene import orbit_camera

W, H, SPP = 40, 24, 5


def test_sample_and_ray_slices_cover_exactly_once():
    for world in (1, 2, 3, 4, 8):
        for total in (0, 1, 5, 8, 256, 1024):
            seen = []
            for r in range(world):
                first, count, stride = multi_gpu.sample_slice(total, r, world)
                seen += [first + k * stride for k in range(count)]
            assert sorted(seen) == list(range(total))
        for n in (0, 1, 7, 64 << 20):
            cuts = [multi_gpu.ray_slice(n, r, world) for r in range(world)]
            assert cuts[0][0] == 0 and cuts[-1][1] == n
            assert all(a[1] == b[0] for a, b in zip(cuts, cuts[1:]))
            assert max(e - b for b, e in cuts) - min(e - b for b, e in cuts) <= 1
    with pytest.raises(ValueError):
        multi_gpu.sample_slice(4, 2, 2)


def _worker(rank, world, port, out_path):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))
    dist.init_process_group("gloo", rank=rank, world_size=world)
    orc = O.Oracle("port")
    sb, camkw, _, bounce = G.small_scene("testscene")
    desc = sb.finalize()
    cam = orbit_camera(W, H, **camkw)
    first, count, stride = multi_gpu.sample_slice(SPP, rank, world)
    acc, _ = orc.render(desc, cam, W, H, first, count, max_bounce=bounce, sample_stride=stride, threads=2)
    t = torch.from_numpy(acc)
    multi_gpu.reduce_accumulator(t, world, dst=0)
    # max-over-ranks timing plumbing used by bench.py
    ms = torch.tensor([float(rank + 1)], dtype=torch.float64)
    dist.all_reduce(ms, op=dist.ReduceOp.MAX)
    if rank == 0:
        np.save(out_path, np.concatenate([t.numpy().reshape(-1), ms.numpy()]))
    dist.barrier()
    dist.destroy_process_group()


def test_two_rank_sample_split_and_reduce(tmp_path):
    if not O.available("port"):
        pytest.fail("oracle/liboracle.so missing: make -C oracle")
    with socket.socket() as s:
        s.bind(("127.0.0.1", 0))
        port = s.getsockname()[1]
    out = str(tmp_path / "acc.npy")
    mp.spawn(_worker, args=(2, port, out), nprocs=2, join=True)
    got = np.load(out)
    assert got[-1] == 2.0
    orc = O.Oracle("port")
    sb, camkw, _, bounce = G.small_scene("testscene")
    whole, _ = orc.render(sb.finalize(), orbit_camera(W, H, **camkw), W, H, 0, SPP, max_bounce=bounce)
    # same set of samples, different order of float additions
    assert np.allclose(got[:-1].reshape(H, W, 4), whole, rtol=1e-5, atol=1e-5)
