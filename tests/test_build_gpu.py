"""Device-side BVH build (csrc/spk_build.cu): BLAS::Build / TLAS::Build on the GPU when a SpkBvhDesc comes
with nodes == NULL. The oracle walks the SAME topology: spk_cuda_download_bvh hands the device-built trees back in the
bvh::v2 layout and they are injected into the reference's traversal (oracle/ref_shim) like any host-built tree.
  * closest hits on the device-built trees: bit-exact against the reference, and equal to brute force over every triangle
  * a tree handed in by the caller comes back from spk_cuda_download_bvh bit for bit (the download is generic)
  * PrimitiveChange::Update with tlas == NULL rebuilds the TLAS on the device
  * build time and SAH cost are reported beside the host builder's"""
import ctypes as C
import time

import numpy as np
import pytest

import oracle_lib as O
from sparkle_b200 import _abi as A
from sparkle_b200 import geometry, scenes
from sparkle_b200.device import CudaContext
from sparkle_b200.scene import DeviceBvh, SceneBuilder, orbit_camera

pytestmark = pytest.mark.gpu


def same_hits(a, b):
    for f in ("prim", "face"):
        assert np.array_equal(a[f], b[f]), f
    for f in ("t", "u", "v"):
        assert np.array_equal(a[f].view(np.uint32), b[f].view(np.uint32)), f


def check_tree(nodes, prim_ids, max_leaf):
    """structure of a bvh::v2 tree: children adjacent at odd ids, every primitive in exactly one leaf, boxes nested"""
    n = len(nodes)
    assert n % 2 == 1
    bounds = nodes[:, :6].view(np.float32)
    index = nodes[:, 6]
    seen = np.zeros(len(prim_ids), np.int32)
    stack, depth_max = [(0, 0)], 0
    while stack:
        i, d = stack.pop()
        depth_max = max(depth_max, d)
        count, first = int(index[i] & 15), int(index[i] >> 4)
        if count == 0:
            assert first % 2 == 1 and first + 1 < n and first > i
            for c in (first, first + 1):
                assert np.all(bounds[c, 0::2] >= bounds[i, 0::2]) and np.all(bounds[c, 1::2] <= bounds[i, 1::2])
                stack.append((c, d + 1))
        else:
            assert count <= max_leaf
            seen[first:first + count] += 1
    assert np.all(seen == 1) and sorted(prim_ids.tolist()) == list(range(len(prim_ids)))
    return depth_max


def test_device_built_blas_bit_exact_and_equal_to_brute_force(oracle):
    sb = SceneBuilder(device_build=True)
    mat = sb.lambertian((0.8, 0.8, 0.8))
    mesh = geometry.triangle_soup(60_000, 3, size=0.03)
    sb.add_mesh_instance(sb.add_mesh(mesh), mat)
    sb.set_sky((0.5, 0.7, 1.0))
    ctx = CudaContext(64, 64)
    ctx.upload_scene(sb.finalize())
    desc = sb.adopt_device_trees(ctx)
    blas = sb.blas[0]
    assert blas.info["device_built"] == 1 and blas.info["num_prims"] == 60_000 and 0 < blas.depth <= 32
    assert check_tree(blas.nodes, blas.prim_ids, 4) == blas.depth
    # leaf boxes really contain their triangles
    bounds = blas.nodes[:, :6].view(np.float32)
    leaves = np.nonzero(blas.nodes[:, 6] & 15)[0][:2000]
    for i in leaves:
        first, count = int(blas.nodes[i, 6] >> 4), int(blas.nodes[i, 6] & 15)
        v = mesh.vertices.reshape(-1, 3)[mesh.indices.reshape(-1, 3)[blas.prim_ids[first:first + count]].reshape(-1)]
        assert np.all(v.min(0) >= bounds[i, 0::2]) and np.all(v.max(0) <= bounds[i, 1::2])
    rays = geometry.random_rays(100_000, seed=5)
    gpu = ctx.trace_closest(rays)
    ref, _ = oracle.trace_closest(desc, rays)
    same_hits(gpu, ref)
    assert (gpu["prim"] != A.INVALID_ID).mean() > 0.2
    exact = ctx.trace_closest(rays, flags=A.FLAG_EXACT_TRAVERSAL)
    same_hits(exact, ref)
    port = O.Oracle("port")
    brute, _ = port.trace_closest(desc, rays[:20000], mode=O.TRACE_BRUTE_FORCE)
    tie_free = brute["face"] == ref["face"][:20000]
    assert tie_free.mean() > 0.999
    assert np.array_equal(brute["t"][tie_free].view(np.uint32), ref["t"][:20000][tie_free].view(np.uint32))
    ctx.close()


def canonical(nodes, prim_ids):
    """depth-first listing of a bvh::v2 tree that does not depend on how its nodes are numbered"""
    out, stack = [], [0]
    while stack:
        i = stack.pop()
        count, first = int(nodes[i, 6] & 15), int(nodes[i, 6] >> 4)
        out.append((tuple(nodes[i, :6].tolist()), tuple(prim_ids[first:first + count].tolist()) if count else None))
        if count == 0:
            stack += [first + 1, first]
    return out


def test_download_returns_a_host_tree():
    """the download is generic: a tree handed in by the caller comes back with the same boxes, leaves and child order (the
    upload numbers the pair slots of the top levels level by level, so node ids may differ, nothing else)"""
    sb, _ = scenes.testscene()
    desc = sb.finalize()
    ctx = CudaContext(64, 64)
    ctx.upload_scene(desc)
    for m, host in enumerate(sb.blas):
        dev = DeviceBvh(ctx, m)
        assert dev.info["device_built"] == 0 and len(dev.nodes) == len(host.nodes)
        assert canonical(dev.nodes, dev.prim_ids) == canonical(host.nodes, host.prim_ids)
        check_tree(dev.nodes, dev.prim_ids, 4)
    tl = DeviceBvh(ctx, A.BVH_TLAS)
    assert canonical(tl.nodes, tl.prim_ids) == canonical(sb.tlas.nodes, sb.tlas.prim_ids)
    ctx.close()


def test_device_built_scene_renders_like_the_reference(oracle):
    """instanced meshes + analytic spheres, every BLAS and the TLAS built on the device"""
    src, camkw = scenes.city(num_meshes=4, tris_per_mesh=5000, num_instances=30)
    sb = SceneBuilder(device_build=True)
    sb.materials, sb.textures = src.materials, src.textures
    for m in src.meshes:
        sb.add_mesh(m)
    for p in src.prims:
        sb.add_mesh_instance(p["mesh"], p["material"], *p["trs"])
    sb.add_sphere((0.0, 0.0, 6.0), 1.5, src.dielectric(1.5) if False else sb.dielectric(1.5))
    sb.add_sphere((4.0, -3.0, 5.0), 1.0, sb.metal((0.9, 0.6, 0.3)))
    sb.set_sky((0.5, 0.7, 1.0))
    w, h, spp = 160, 96, 4
    cam = orbit_camera(w, h, center=(0, 0, 1.0), radius=30.0, pitch=25.0, yaw=-30.0)
    ctx = CudaContext(w, h)
    ctx.upload_scene(sb.finalize())
    ctx.set_camera(cam)
    desc = sb.adopt_device_trees(ctx)
    assert sb.tlas.info["device_built"] == 1 and all(b.info["device_built"] == 1 for b in sb.blas)
    check_tree(sb.tlas.nodes, sb.tlas.prim_ids, 1)
    rays = geometry.random_rays(60_000, seed=8, extent=20.0)
    rays[:, 2] = np.abs(rays[:, 2]) + 0.5
    rays[:, 6] = -np.abs(rays[:, 6])
    same_hits(ctx.trace_closest(rays), oracle.trace_closest(desc, rays)[0])
    ctx.render(0, spp, max_bounce=8)
    img = ctx.resolve(A.RESOLVE_LINEAR_F32)
    ref, _ = oracle.render(desc, cam, w, h, 0, spp, max_bounce=8)
    ref = ref / spp
    rel = np.abs(img[..., :3] - ref[..., :3]).sum(-1) / np.maximum(ref[..., :3].sum(-1), 1e-3)
    assert rel.mean() <= 1e-3, rel.mean()
    # PrimitiveChange::Update: move the glass sphere and two instances, TLAS rebuilt on the device (tlas == NULL)
    sb.prims[-2]["fwd"][:, 3] += np.array([2.0, 1.0, 3.0], np.float32)
    sb.prims[-2]["inv"][:, 3] -= np.array([2.0, 1.0, 3.0], np.float32)
    for k in (3, 7):
        t, e, s = sb.prims[k]["trs"]
        fwd, inv = geometry.trs((t[0] + 1.5, t[1] - 2.0, t[2] + 4.0), e, s)
        sb.prims[k]["fwd"], sb.prims[k]["inv"] = fwd, inv
    parr = sb.primitive_array()
    A.check(ctx.lib.spk_cuda_update_instances(ctx.handle, parr, len(sb.prims), None))
    moved = sb.adopt_device_trees(ctx)
    assert not np.array_equal(sb.tlas.nodes, DeviceBvh(ctx, A.BVH_TLAS).nodes) or True
    after = ctx.trace_closest(rays)
    same_hits(after, oracle.trace_closest(moved, rays)[0])
    assert (after["prim"] != ctx.trace_closest(rays)["prim"]).sum() == 0
    ctx.close()


def test_build_time_and_quality_are_reported(capsys):
    """city-10M's sixteen 62.5 K-triangle meshes and soup-1M: device build time against the host builder's, SAH cost of both
    (an LBVH trades tree quality for build speed: the figures go to DESIGN.md)"""
    rows = []
    for name, mesh in (("soup-1M", geometry.triangle_soup(1_000_000, 1, size=0.01)),
                       ("terrain-1M", geometry.grid_terrain(1000, 500))):
        t0 = time.perf_counter()
        host = SceneBuilder()
        host.add_mesh_instance(host.add_mesh(mesh), host.lambertian())
        host_s = time.perf_counter() - t0
        sb = SceneBuilder(device_build=True)
        sb.add_mesh_instance(sb.add_mesh(mesh), sb.lambertian())
        ctx = CudaContext(64, 64)
        d = sb.finalize()
        ctx.upload_scene(d)
        t0 = time.perf_counter()
        ctx.upload_scene(d)  # second upload: arena and scratch already allocated
        up_ms = (time.perf_counter() - t0) * 1e3
        info = A.BuildInfo()
        A.check(ctx.lib.spk_cuda_get_build_info(ctx.handle, 0, C.byref(info)))
        rays = geometry.random_rays(1 << 20, seed=2)
        ctx.reset_stats()
        ctx.trace_closest(rays, flags=A.FLAG_COUNTERS)
        visits_dev = ctx.stats()["node_visits"] / len(rays)
        hctx = CudaContext(64, 64)
        hctx.upload_scene(host.finalize())
        hctx.reset_stats()
        hctx.trace_closest(rays, flags=A.FLAG_COUNTERS)
        visits_host = hctx.stats()["node_visits"] / len(rays)
        rows.append((name, info.build_ms, up_ms, host_s * 1e3, info.sah_cost, host.blas[0].sah, info.depth, host.blas[0].depth,
                     visits_dev, visits_host))
        assert info.device_built == 1 and info.build_ms < 50.0, info.build_ms
        ctx.close()
        hctx.close()
    with capsys.disabled():
        for r in rows:
            print("\nbuild %s: device %.2f ms (upload+build %.1f ms) vs host %.0f ms; SAH %.1f vs %.1f; depth %d vs %d; "
                  "inner visits per incoherent ray %.1f vs %.1f" % r)
