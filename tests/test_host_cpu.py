"""CPU tests of the host side (no GPU): the BVH builder's invariants, the `pipeline` switch (Str2Enum semantics of
core/Enum.h:20-32 and the new `cuda` enumerator), the no-fallback rule of Renderer::CreateRenderer, and the C++
TestScene recipe against the Python one."""
import numpy as np
import pytest

from sparkle_b200 import _abi as A
from sparkle_b200 import app, geometry, scenes
from sparkle_b200.scene import HostBvh


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
