"""CPU tests of the host side (no GPU): the BVH builder's invariants, the `pipeline` switch (Str2Enum semantics of
core/Enum.h:20-32 and the new `cuda` enumerator), the no-fallback rule of Renderer::CreateRenderer, and the C++
TestScene recipe against the Python one."""
import os

import numpy as np
import pytest

import golden_cases as G
import oracle_lib as O
from sparkle_b200 import _abi as A
from sparkle_b200 import app, geometry, scenes
from sparkle_b200.scene import HostBvh, orbit_camera


GOLDEN = os.path.join(os.path.dirname(os.path.abspath(__file__)), "golden", "reference_vectors.npz")


def host_trs(t, e, s):
    return geometry.trs(t, e, s)


def host_node(t, q):
    t, q = np.ascontiguousarray(t, np.float32), np.ascontiguousarray(q, np.float32)
    fwd, inv = np.empty(12, np.float32), np.empty(12, np.float32)
    A.host_lib().spk_host_gltf_node(*[a.ctypes.data_as(A.fp) for a in (t, q, fwd, inv)])
    return fwd.reshape(3, 4), inv.reshape(3, 4)


def test_host_math_reproduces_the_reference_bit_for_bit():
    """a17 + transforms: the product's Transform / glTF-node / orbit-camera math (host/ref_math.cpp, used by both the C++
    host and these Python bindings) against the reference's OWN Transform::UpdateTransformData (Transform.cpp:52-64) and
    CameraRenderProxy::OnTransformDirty (CameraRenderProxy.cpp:113-136), via the committed golden vectors
    (hostmath_*, generated from oracle/_ref by tests/golden/make_golden.py) and, where the reference library is
    built, live on a larger seeded set. Same float32 VALUES everywhere (-0 == +0)."""
    golden = dict(np.load(GOLDEN))
    got = G.evaluate_host_math(host_trs, host_node, orbit_camera)
    for k, v in got.items():
        assert v.dtype == np.float32 and np.array_equal(v, golden[k]), f"{k}: {int((v != golden[k]).sum())} values differ"
    if not O.available("reference"):
        return
    ref = O.Oracle("reference")
    g = np.random.default_rng(2024)
    for _ in range(3000):
        t, e, s = g.uniform(-50, 50, 3), g.uniform(-6.3, 6.3, 3), g.uniform(0.05, 20, 3)
        a, b = host_trs(t, e, s), ref.kat_transform(t, e, s)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (t, e, s)
        q = g.normal(size=4)
        a, b = host_node(t, q), ref.kat_gltf_node(t, q)
        assert np.array_equal(a[0], b[0]) and np.array_equal(a[1], b[1]), (t, q)
        kw = dict(center=tuple(g.uniform(-9, 9, 3)), radius=float(g.uniform(0.5, 90)), pitch=float(g.uniform(-89, 89)),
                  yaw=float(g.uniform(-180, 180)), focal_length=float(g.uniform(0.01, 0.1)),
                  aperture=float(g.uniform(0.95, 22)))
        w, h = int(g.integers(2, 4096)), int(g.integers(2, 2304))
        assert np.array_equal(G.camera_floats(orbit_camera(w, h, **kw)), G.camera_floats(ref.kat_camera(w, h, **kw))), kw


def _has_gpu():
    try:
        import torch
        return torch.cuda.is_available()
    except Exception:
        return False


def check_tree(bvh, boxes, max_leaf):
    nodes = bvh.nodes
    bounds = nodes[:, :6].view(np.float32)
    index = nodes[:, 6]
    seen = np.zeros(len(boxes), np.int32)
    stack = [(0, 0)]
    depth = 0
    while stack:
        n, d = stack.pop()
        depth = max(depth, d)
        first, count = index[n] >> 4, index[n] & 15
        b = bounds[n]
        if count == 0:
            assert first % 2 == 1 and first + 1 < len(nodes)  # children adjacent, pairs at odd/even ids
            for c in (first, first + 1):
                cb = bounds[c]
                assert (cb[0::2] >= b[0::2]).all() and (cb[1::2] <= b[1::2]).all(), "child box escapes its parent"
                stack.append((c, d + 1))
        else:
            assert count <= max_leaf
            ids = bvh.prim_ids[first:first + count]
            seen[ids] += 1
            pb = boxes[ids]
            assert (pb[:, :3] >= b[0::2]).all() and (pb[:, 3:] <= b[1::2]).all(), "primitive outside its leaf box"
    assert (seen == 1).all(), "every primitive in exactly one leaf"
    assert depth <= 32  # SmallStack<Index, 32> of the reference's traversal
    return depth


@pytest.mark.parametrize("n,max_leaf", [(1, 4), (2, 1), (7, 4), (1000, 4), (5000, 8)])
def test_bvh_builder_invariants(n, max_leaf):
    g = np.random.default_rng(n)
    c = g.uniform(-10, 10, (n, 3)).astype(np.float32)
    h = g.uniform(0.01, 0.5, (n, 3)).astype(np.float32)
    boxes = np.concatenate([c - h, c + h], axis=1)
    bvh = HostBvh.from_boxes(boxes, max_leaf=max_leaf)
    assert len(bvh.prim_ids) == n and sorted(bvh.prim_ids.tolist()) == list(range(n))
    depth = check_tree(bvh, boxes, max_leaf)
    assert depth == bvh.depth or n == 1


def test_bvh_builder_degenerate_and_errors():
    same = np.tile(np.array([[0, 0, 0, 1, 1, 1]], np.float32), (100, 1))  # identical centroids: median fallback
    bvh = HostBvh.from_boxes(same, max_leaf=2)
    check_tree(bvh, same, 2)
    with pytest.raises(RuntimeError):
        HostBvh.from_boxes(np.zeros((0, 6), np.float32))
    with pytest.raises(RuntimeError):
        HostBvh.from_boxes(same, max_leaf=16)  # prim_count is a 4-bit field
    mesh = geometry.triangle_soup(300, seed=3, size=0.1)
    b = HostBvh.from_mesh(mesh)
    v = mesh.vertices[mesh.indices.reshape(-1, 3)]
    boxes = np.concatenate([v.min(1), v.max(1)], axis=1)
    check_tree(b, boxes, 4)


def test_pipeline_switch():
    # Str2Enum is case-insensitive (Enum.h:20-32); `cuda` parses beside the reference's four pipelines
    for name, value in app.PIPELINES.items():
        assert app.parse_pipeline(name) == value
        assert app.parse_pipeline(name.upper()) == value
        assert app.parse_pipeline(name.capitalize()) == value
    assert app.parse_pipeline("vulkan") == -1 and app.parse_pipeline("") == -1


def test_create_renderer_has_no_fallback():
    lib = A.host_lib()
    # the reference's own pipelines are not provided by this library
    with pytest.raises(RuntimeError, match="cuda"):
        app.HostApp("--pipeline forward --width 64 --height 48")
    with pytest.raises(RuntimeError):
        app.HostApp("--pipeline cuda --width")             # odd token count
    with pytest.raises(RuntimeError, match="unknown config"):
        app.HostApp("--pipeline cuda --no_such_cvar 1")
    with pytest.raises(RuntimeError, match="bad value"):
        app.HostApp("--pipeline cuda --bounce many")
    if not _has_gpu():
        with pytest.raises(RuntimeError, match="(?i)device"):
            app.HostApp("--pipeline cuda --width 64 --height 48 --max_spp 2")
    assert lib.spk_host_last_error() is not None


def test_cpp_testscene_matches_python_recipe():
    """both host sides place the ten fixed-seed-42 spheres identically (SceneManager.cpp:51-105)"""
    sb, _ = scenes.testscene(models=False)
    spheres = [p for p in sb.prims if p["kind"] == A.PRIM_SPHERE]
    assert len(spheres) == 13
    # the C++ scene is only reachable through an app, which needs a device: compare via the python mirror of the
    # recipe's invariants here (no overlaps, inside the 13 m disk, radius range), the GPU test compares bit for bit
    for p in spheres[3:]:
        c = (p["lo"] + p["hi"]) * 0.5
        r = (p["hi"][2] - p["lo"][2]) * 0.5
        assert 0.4 <= r <= 1.0 and abs(c[2] - r) < 1e-6 and np.hypot(c[0], c[1]) <= 13.0 + 1e-4
    for i, a in enumerate(spheres):
        for b in spheres[i + 1:]:
            ca, cb = (a["lo"] + a["hi"]) * 0.5, (b["lo"] + b["hi"]) * 0.5
            ha, hb = (a["hi"] - a["lo"]) * 0.5, (b["hi"] - b["lo"]) * 0.5
            assert not np.all(np.abs(ca - cb) < ha + hb)
