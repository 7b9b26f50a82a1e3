"""Host-side geometry helpers for building inputs of the `cuda` pipeline: the per-(pixel,sample) RNG of the
reference (for seeded procedural content), mesh generators with the attribute set io/Mesh.h requires
(vertices, normals, tangents, uvs, indices) and TRS matrices laid out like Transform::GetMatrix().

Reference: libraries/source/io/Mesh.cpp:7-200 (unit sphere / unit cube), core/math/Sampler.h:32-53
(RandomUnit<true>, UnitDisk<true>), thirdparty/xoshiro_cpp/XoshiroCpp.hpp (Xoshiro128Plus, SplitMix64)."""
import math

import numpy as np

f32 = np.float32
M32 = 0xFFFFFFFF
M64 = 0xFFFFFFFFFFFFFFFF


class Xoshiro128Plus:
    """XoshiroCpp::Xoshiro128Plus(seed): the seed is zero-extended to 64 bits and each of the four state words is a
    SplitMix64 output truncated to its low 32 bits (XoshiroCpp.hpp:684-690,1305-1330). KAT: seed 0 ->
    7b1dcdaf a1b965f4 8009454f 724c81ec (tests/test_host.py)."""

    def __init__(self, seed):
        x = seed & M64
        s = []
        for _ in range(4):
            x = (x + 0x9E3779B97F4A7C15) & M64
            z = x
            z = ((z ^ (z >> 30)) * 0xBF58476D1CE4E5B9) & M64
            z = ((z ^ (z >> 27)) * 0x94D049BB133111EB) & M64
            z = z ^ (z >> 31)
            s.append(z & M32)
        self.s = s

    def next(self):
        s = self.s
        r = (s[0] + s[3]) & M32
        t = (s[1] << 9) & M32
        s[2] ^= s[0]
        s[3] ^= s[1]
        s[1] ^= s[2]
        s[0] ^= s[3]
        s[2] ^= t
        s[3] = ((s[3] << 11) | (s[3] >> 21)) & M32
        return r

    def unit(self):
        """sampler::RandomUnit: float(r) / float(UINT32_MAX) (both int->float conversions round to nearest)"""
        return f32(f32(self.next()) / f32(4294967295))

    def unit_disk(self):
        """sampler::UnitDisk (Sampler.h:43-53), float32 steps"""
        rand = self.unit()
        theta = f32(f32(self.unit() * f32(2.0)) * f32(math.pi))
        r = f32(np.sqrt(rand))
        return f32(r * f32(np.cos(theta))), f32(r * f32(np.sin(theta)))


def trs(translate=(0, 0, 0), euler=(0, 0, 0), scale=(1, 1, 1)):
    """Transform(translate, euler, scale): Translation * (Rx * Ry * Rz as quaternions) * Scaling and its affine inverse
    (Transform.cpp:52-64, Utilities.h:392-396), computed by the host library in the reference's float operation order
    (sparkle_b200/host/ref_math.cpp; bit-identical to the reference's Eigen code, tests/test_host_cpu.py).
    Returns (fwd, inv) as float32 3x4 row-major arrays."""
    from . import _abi as A
    t, e, s = (np.ascontiguousarray(v, f32) for v in (translate, euler, scale))
    fwd, inv = np.empty(12, f32), np.empty(12, f32)
    A.host_lib().spk_host_trs(t.ctypes.data_as(A.fp), e.ctypes.data_as(A.fp), s.ctypes.data_as(A.fp),
                              fwd.ctypes.data_as(A.fp), inv.ctypes.data_as(A.fp))
    return fwd.reshape(3, 4), inv.reshape(3, 4)


class MeshData:
    def __init__(self, vertices, normals, tangents, uvs, indices, name="mesh"):
        self.vertices = np.ascontiguousarray(vertices, f32).reshape(-1, 3)
        self.normals = np.ascontiguousarray(normals, f32).reshape(-1, 3)
        self.tangents = np.ascontiguousarray(tangents, f32).reshape(-1, 4)
        self.uvs = None if uvs is None else np.ascontiguousarray(uvs, f32).reshape(-1, 2)
        self.indices = np.ascontiguousarray(indices, np.uint32).reshape(-1)
        self.name = name
        assert len(self.vertices) == len(self.normals) == len(self.tangents)
        assert self.indices.size % 3 == 0 and self.indices.size > 0
        assert self.indices.max() < len(self.vertices)

    @property
    def num_faces(self):
        return self.indices.size // 3

    def bounds(self):
        return self.vertices.min(axis=0), self.vertices.max(axis=0)


def _major_axis(n):
    """utilities::GetPossibleMajorAxis, vectorised"""
    inv_sqrt3 = f32(1.0 / math.sqrt(3.0))
    out = np.zeros_like(n)
    a = np.abs(n[:, 0]) < inv_sqrt3
    b = ~a & (np.abs(n[:, 1]) < inv_sqrt3)
    c = ~a & ~b
    out[a, 0] = 1
    out[b, 1] = 1
    out[c, 2] = 1
    return out


def tangents_from_normals(normals):
    m = _major_axis(normals)
    t = np.cross(np.cross(normals, m), normals).astype(f32)
    return np.concatenate([t, np.ones((len(t), 1), f32)], axis=1)


def unit_cube():
    """Mesh::GetUnitCube: 6 quads on [-1,1]^3, 24 vertices, per-face normal/tangent, uv per corner"""
    faces = [  # normal, tangent(u direction)
        ((0, 0, 1), (1, 0, 0)), ((0, 0, -1), (-1, 0, 0)), ((1, 0, 0), (0, 1, 0)), ((-1, 0, 0), (0, -1, 0)),
        ((0, 1, 0), (-1, 0, 0)), ((0, -1, 0), (1, 0, 0))]
    V, N, T, UV, I = [], [], [], [], []
    for n, t in faces:
        n = np.array(n, np.float64)
        t = np.array(t, np.float64)
        b = np.cross(n, t)
        corners = [n - t - b, n + t - b, n + t + b, n - t + b]  # BL, BR, TR, TL
        base = len(V)
        V += corners
        N += [n] * 4
        T += [np.append(t, 1.0)] * 4
        UV += [(0, 1), (1, 1), (1, 0), (0, 0)]
        I += [base, base + 2, base + 1, base, base + 3, base + 2]
    # winding: make every triangle's geometric normal agree with the face normal
    V = np.array(V)
    I = np.array(I).reshape(-1, 3)
    N = np.array(N)
    gn = np.cross(V[I[:, 1]] - V[I[:, 0]], V[I[:, 2]] - V[I[:, 0]])
    flip = np.einsum("ij,ij->i", gn, N[I[:, 0]]) < 0
    I[flip] = I[flip][:, [0, 2, 1]]
    return MeshData(V, N, np.array(T), np.array(UV), I, "unit_cube")


def uv_sphere(v_size=32, radius=1.0, name="uv_sphere"):
    """Mesh::GetUnitSphere layout (Mesh.cpp:7-86): (2v+1)x(v+1) vertices, 2v*v quads"""
    u_size = v_size * 2
    u = np.arange(u_size + 1)
    v = np.arange(v_size + 1)
    theta = (2 * math.pi * u / u_size + math.pi)[None, :]
    phi = (math.pi * v / v_size)[:, None]
    x = np.cos(theta) * np.sin(phi)
    y = -np.cos(phi) * np.ones_like(theta)
    z = np.sin(theta) * np.sin(phi)
    P = np.stack([x, y, z], -1).reshape(-1, 3)
    P /= np.linalg.norm(P, axis=1, keepdims=True)
    normals = P.astype(f32)
    idx = []
    for yy in range(v_size):
        for xx in range(u_size):
            vi = yy * (u_size + 1) + xx
            idx += [vi, vi + u_size + 1, vi + 1, vi + 1, vi + u_size + 1, vi + u_size + 2]
    uvs = np.stack([np.tile(u / u_size, v_size + 1), np.repeat(v / v_size, u_size + 1)], -1)
    I = np.array(idx).reshape(-1, 3)
    # drop the degenerate pole triangles (zero area: |a| < Eps never hits, they only cost memory)
    Pv = P * radius
    area = np.linalg.norm(np.cross(Pv[I[:, 1]] - Pv[I[:, 0]], Pv[I[:, 2]] - Pv[I[:, 0]]), axis=1)
    I = I[area > 1e-12]
    return MeshData(Pv, normals, tangents_from_normals(normals), uvs, I, name)


def grid_terrain(nx, ny, size=2.0, height=0.15, seed=7, name="terrain"):
    """(nx x ny) quads -> 2*nx*ny triangles; heightfield = sum of seeded sinusoids (value-noise stand-in)"""
    rng = np.random.default_rng(seed)
    xs = np.linspace(-size / 2, size / 2, nx + 1)
    ys = np.linspace(-size / 2, size / 2, ny + 1)
    X, Y = np.meshgrid(xs, ys)
    Z = np.zeros_like(X)
    for _ in range(8):
        fx, fy = rng.uniform(1, 12, 2)
        ph = rng.uniform(0, 2 * math.pi, 2)
        Z += rng.uniform(0.2, 1.0) / (fx + fy) * np.sin(fx * X * math.pi + ph[0]) * np.cos(fy * Y * math.pi + ph[1])
    Z *= height / max(1e-9, np.abs(Z).max())
    P = np.stack([X, Y, Z], -1).reshape(-1, 3)
    gy, gx = np.gradient(Z, ys, xs)
    Nn = np.stack([-gx, -gy, np.ones_like(Z)], -1).reshape(-1, 3)
    Nn /= np.linalg.norm(Nn, axis=1, keepdims=True)
    Nn = Nn.astype(f32)
    uv = np.stack([(X - xs[0]) / size, (Y - ys[0]) / size], -1).reshape(-1, 2)
    j, i = np.meshgrid(np.arange(ny), np.arange(nx), indexing="ij")
    v00 = (j * (nx + 1) + i).reshape(-1)
    v10 = v00 + 1
    v01 = v00 + nx + 1
    v11 = v01 + 1
    I = np.stack([v00, v10, v11, v00, v11, v01], -1).reshape(-1, 3)
    return MeshData(P, Nn, tangents_from_normals(Nn), uv, I, name)


def triangle_soup(num_tris, seed=1, extent=1.0, size=0.01, name="soup"):
    """SURVEY 8d C3 'soup': centres uniform in [-extent,extent]^3, three offsets uniform in [-size,size]^3"""
    rng = np.random.default_rng(seed)
    c = rng.uniform(-extent, extent, (num_tris, 1, 3))
    off = rng.uniform(-size, size, (num_tris, 3, 3))
    P = (c + off).astype(f32).reshape(-1, 3)
    I = np.arange(num_tris * 3, dtype=np.uint32)
    e1 = P[1::3] - P[0::3]
    e2 = P[2::3] - P[0::3]
    n = np.cross(e1, e2)
    n /= np.maximum(np.linalg.norm(n, axis=1, keepdims=True), 1e-20)
    Nn = np.repeat(n.astype(f32), 3, axis=0)
    uv = np.tile(np.array([[0, 0], [1, 0], [0, 1]], f32), (num_tris, 1))
    return MeshData(P, Nn, tangents_from_normals(Nn), uv, I, name)


def random_rays(n, seed=2, extent=1.5):
    """SURVEY 8d C3 rays: origin uniform in [-extent,extent]^3, direction uniform on S^2 (normalised in float32 the
    way Vector3::normalized does: a / sqrt(x*x + (y*y + z*z)))."""
    rng = np.random.default_rng(seed)
    rays = np.zeros((n, 8), f32)
    rays[:, 0:3] = rng.uniform(-extent, extent, (n, 3)).astype(f32)
    d = rng.standard_normal((n, 3)).astype(f32)
    nrm = np.sqrt(d[:, 0] * d[:, 0] + (d[:, 1] * d[:, 1] + d[:, 2] * d[:, 2]))
    rays[:, 4:7] = d / nrm[:, None]
    return rays
