"""Leaf size of the device LBVH (SPK_OPT_BUILD_MAX_LEAF) against tracing speed: soup-1M and terrain-1M, 16 Mi incoherent
rays (and the host SAH tree for reference)."""
import os, sys
sys.path.insert(0, os.path.join(os.path.dirname(os.path.abspath(__file__)), ".."))
import torch
from sparkle_b200 import _abi as A, geometry
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import SceneBuilder
n = 16 << 20
rays = torch.from_numpy(geometry.random_rays(n, seed=2)).cuda()
hits = torch.empty((n, 5), dtype=torch.float32, device="cuda")
for name, mesh in (("soup-1M", geometry.triangle_soup(1_000_000, 1, size=0.01)), ("terrain-1M", geometry.grid_terrain(1000, 500))):
    for leaf in (0, 1, 2, 3, 4, 8):
        sb = SceneBuilder(device_build=leaf > 0)
        sb.add_mesh_instance(sb.add_mesh(mesh), sb.lambertian())
        ctx = CudaContext(64, 64)
        if leaf:
            ctx.set_option(A.OPT_BUILD_MAX_LEAF, leaf)
        ctx.upload_scene(sb.finalize())
        best = 1e30
        for rep in range(3):
            ctx.trace_closest_device(rays.data_ptr(), n, hits.data_ptr(), 0); ctx.synchronize()
            if rep: best = min(best, ctx.stats()["last_ms"])
        ctx.reset_stats(); ctx.trace_closest_device(rays.data_ptr(), 1 << 20, hits.data_ptr(), A.FLAG_COUNTERS); st = ctx.stats()
        print(f"{name} {'host SAH' if not leaf else 'LBVH max_leaf %d' % leaf}: {best:.2f} ms = {n/best/1e3:.0f} Mrays/s; "
              f"visits/ray {st['node_visits']/(1<<20):.1f} tri/ray {st['tri_tests']/(1<<20):.2f}", flush=True)
        ctx.close()
