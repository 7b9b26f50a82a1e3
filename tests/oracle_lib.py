"""TEST INFRASTRUCTURE: ctypes loader for the CPU oracle (oracle/oracle.h). Only tests/, __graft_entry__.smoke() and
bench.py's cpu_baseline / --impl reference legs may import this. kind = "reference" (oracle/_ref/libsparkle_ref.so, the
reference's own CPU path tracer compiled in place) or "port" (oracle/liboracle.so, the plain C++ restatement)."""
import ctypes as C
import os

import numpy as np

from sparkle_b200 import _abi as A
from sparkle_b200.device import HIT_DTYPE

ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
PATHS = {"reference": os.path.join(ROOT, "oracle", "_ref", "libsparkle_ref.so"),
         "port": os.path.join(ROOT, "oracle", "liboracle.so")}
TRACE_BVH_REFERENCE, TRACE_BRUTE_FORCE, TRACE_BVH_CULLED = 0, 1, 2


class OrcCounters(C.Structure):
    _fields_ = [(n, C.c_uint64) for n in ("paths", "rays", "node_visits", "tri_tests", "inst_visits", "sphere_tests",
                                          "rng_draws")]


_cache = {}


def available(kind):
    return os.path.exists(PATHS[kind])


def load(kind="reference"):
    if kind in _cache:
        return _cache[kind]
    lib = C.CDLL(PATHS[kind])
    lib.orc_kind.restype = C.c_char_p
    f32, u32, fp = C.c_float, C.c_uint32, A.fp
    lib.orc_rng_state.argtypes = [u32, A.u32p]
    lib.orc_rng_units.argtypes = [u32, u32, fp]
    lib.orc_kat_vec.argtypes = [u32, fp, fp, fp, fp]
    lib.orc_kat_libm.argtypes = [u32, f32, f32]
    lib.orc_kat_libm.restype = f32
    lib.orc_kat_unit_disk.argtypes = [u32, fp]
    lib.orc_kat_texture.argtypes = [C.POINTER(A.TextureDesc), f32, f32, fp]
    lib.orc_kat_sky.argtypes = [C.POINTER(A.SkyDesc), fp, fp]
    lib.orc_kat_sample_surface.argtypes = [C.POINTER(A.SceneDesc), u32, u32, fp, fp, fp, fp, fp, fp]
    lib.orc_kat_sample_surface.restype = u32
    lib.orc_trace_closest.argtypes = [C.POINTER(A.SceneDesc), C.c_void_p, C.c_uint64, C.c_void_p, u32, u32,
                                      C.POINTER(OrcCounters)]
    lib.orc_intersect_full.argtypes = [C.POINTER(A.SceneDesc), C.c_void_p, fp]
    lib.orc_intersect_full.restype = u32
    lib.orc_render.argtypes = [C.POINTER(A.SceneDesc), C.POINTER(A.Camera), u32, u32, C.POINTER(A.RenderParams), fp,
                               A.u32p, u32, u32, C.POINTER(OrcCounters)]
    lib.orc_sample_pixel.argtypes = [C.POINTER(A.SceneDesc), C.POINTER(A.Camera), u32, u32, u32, u32, u32, u32, f32,
                                     u32, fp, C.POINTER(OrcCounters)]
    lib.orc_resolve.argtypes = [fp, u32, u32, u32, u32, u32, f32, C.c_void_p]
    if kind == "reference":  # KATs of the reference's scene-graph math (pin the product's host/ref_math.cpp)
        lib.orc_kat_transform.argtypes = [fp, fp, fp, fp, fp]
        lib.orc_kat_gltf_node.argtypes = [fp, fp, fp, fp]
        lib.orc_kat_camera.argtypes = [fp, f32, f32, f32, f32, f32, f32, f32, u32, u32, C.POINTER(A.Camera)]
        lib.orc_kat_decode_hdr.argtypes = [C.c_char_p, u32, C.POINTER(u32), C.POINTER(u32), C.c_void_p, C.c_uint64]
        lib.orc_kat_cook_sky_cube.argtypes = [C.c_void_p, u32, u32, u32, C.c_void_p]
    else:
        lib.orc_set_libm.argtypes = [u32]
    _cache[kind] = lib
    return lib


class Oracle:
    def __init__(self, kind="reference"):
        self.lib = load(kind)
        self.kind = self.lib.orc_kind().decode()

    def trace_closest(self, desc, rays, mode=TRACE_BVH_REFERENCE, threads=8):
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.empty(len(rays), HIT_DTYPE)
        cnt = OrcCounters()
        rc = self.lib.orc_trace_closest(C.byref(desc), rays.ctypes.data, len(rays), hits.ctypes.data, mode, threads,
                                        C.byref(cnt))
        if rc != 0:
            raise RuntimeError(f"orc_trace_closest -> {rc}")
        return hits, cnt

    def render(self, desc, cam, width, height, first_sample, sample_count, max_bounce=8, sample_stride=1,
               accum_mode=A.ACCUM_SUM, accum=None, cumulated=0, threads=8, flags=0, cumulated_cap=0):
        if accum is None:
            accum = np.zeros((height, width, 4), np.float32)
        p = A.RenderParams(first_sample, sample_count, sample_stride, max_bounce, 6.0, accum_mode, flags, cumulated_cap)
        cum = C.c_uint32(cumulated)
        cnt = OrcCounters()
        rc = self.lib.orc_render(C.byref(desc), C.byref(cam), width, height, C.byref(p), accum.ctypes.data_as(A.fp),
                                 C.byref(cum), 0, threads, C.byref(cnt))
        if rc != 0:
            raise RuntimeError(f"orc_render -> {rc}")
        return accum, cum.value

    def sample_pixel(self, desc, cam, width, height, i, j, frame_seed, max_bounce=8):
        out = np.zeros(8, np.float32)
        rc = self.lib.orc_sample_pixel(C.byref(desc), C.byref(cam), width, height, i, j, frame_seed, max_bounce, 6.0,
                                       0, out.ctypes.data_as(A.fp), None)
        if rc != 0:
            raise RuntimeError(f"orc_sample_pixel -> {rc}")
        return out

    def set_libm(self, device_libm):
        """port only: False = glibc float sin/cos/pow like the reference (default); True = what the CUDA path computes (the
        same, except Schlick's pow(x, 5.f) as an exact product rounded once). Process-wide: reset it in a finally block."""
        assert self.kind == "port"
        self.lib.orc_set_libm(1 if device_libm else 0)

    def kat_transform(self, translate, euler, scale):
        t, e, s = (np.ascontiguousarray(v, np.float32) for v in (translate, euler, scale))
        fwd, inv = np.zeros(12, np.float32), np.zeros(12, np.float32)
        self.lib.orc_kat_transform(*[a.ctypes.data_as(A.fp) for a in (t, e, s, fwd, inv)])
        return fwd.reshape(3, 4), inv.reshape(3, 4)

    def kat_gltf_node(self, translation, rotation_xyzw):
        t, q = np.ascontiguousarray(translation, np.float32), np.ascontiguousarray(rotation_xyzw, np.float32)
        fwd, inv = np.zeros(12, np.float32), np.zeros(12, np.float32)
        self.lib.orc_kat_gltf_node(*[a.ctypes.data_as(A.fp) for a in (t, q, fwd, inv)])
        return fwd.reshape(3, 4), inv.reshape(3, 4)

    def kat_decode_hdr(self, data):
        """Image2D::LoadFromFile's HDR branch on a file image: (h, w, 4) uint16 half bits, or None when stb refuses it"""
        w, h = A.u32(0), A.u32(0)
        buf = np.zeros(1 << 24, np.uint16)
        rc = self.lib.orc_kat_decode_hdr(data, len(data), C.byref(w), C.byref(h), buf.ctypes.data, buf.size)
        if rc != 0:
            return None
        return buf[: w.value * h.value * 4].reshape(h.value, w.value, 4).copy()

    def kat_cook_sky_cube(self, equirect_half, size):
        """SkyLightCookJob::Execute's loop: (6, size, size, 4) uint16 half bits"""
        eq = np.ascontiguousarray(equirect_half, np.uint16)
        faces = np.zeros((6, size, size, 4), np.uint16)
        self.lib.orc_kat_cook_sky_cube(eq.ctypes.data, eq.shape[1], eq.shape[0], size, faces.ctypes.data)
        return faces

    def kat_camera(self, width, height, center=(0, 0, 0), radius=25.0, pitch=10.0, yaw=-20.0, focal_length=0.035,
                   sensor_height=0.024, aperture=22.0, exposure=1.0):
        cam = A.Camera()
        c = np.ascontiguousarray(center, np.float32)
        self.lib.orc_kat_camera(c.ctypes.data_as(A.fp), radius, pitch, yaw, focal_length, sensor_height, aperture,
                                exposure, width, height, C.byref(cam))
        return cam

    def resolve(self, accum, fmt, total_samples, exposure=1.0, accum_mode=A.ACCUM_SUM):
        h, w = accum.shape[:2]
        out = np.empty((h, w, 4), {A.RESOLVE_LINEAR_F32: np.float32, A.RESOLVE_TONEMAP_F16: np.float16}.get(fmt, np.uint8))
        rc = self.lib.orc_resolve(np.ascontiguousarray(accum).ctypes.data_as(A.fp), w, h, accum_mode, fmt,
                                  total_samples, exposure, out.ctypes.data)
        if rc != 0:
            raise RuntimeError(f"orc_resolve -> {rc}")
        return out

    # ---- known-answer probes (oracle.h orc_kat_*) ----
    def rng_state(self, seed):
        out = np.zeros(4, np.uint32)
        self.lib.orc_rng_state(seed, out.ctypes.data_as(A.u32p))
        return out

    def rng_units(self, seed, n):
        out = np.zeros(n, np.float32)
        self.lib.orc_rng_units(seed, n, out.ctypes.data_as(A.fp))
        return out

    def kat_vec(self, op, a, b=None, m=None):
        a = np.ascontiguousarray(a, np.float32)
        b = np.ascontiguousarray(b if b is not None else np.zeros(3), np.float32)
        m = np.ascontiguousarray(m if m is not None else np.eye(4)[:3], np.float32).reshape(12)
        out = np.zeros(3, np.float32)
        self.lib.orc_kat_vec(op, a.ctypes.data_as(A.fp), b.ctypes.data_as(A.fp), m.ctypes.data_as(A.fp),
                             out.ctypes.data_as(A.fp))
        return out

    def kat_libm(self, op, x, y=0.0):
        return np.float32(self.lib.orc_kat_libm(op, float(x), float(y)))

    def unit_disk(self, seed):
        out = np.zeros(2, np.float32)
        self.lib.orc_kat_unit_disk(seed, out.ctypes.data_as(A.fp))
        return out

    def texture(self, pixels, fmt, u, v):
        px = np.ascontiguousarray(pixels)
        t = A.TextureDesc(px.ctypes.data, px.shape[1], px.shape[0], fmt, 0)
        out = np.zeros(3, np.float32)
        self.lib.orc_kat_texture(C.byref(t), float(u), float(v), out.ctypes.data_as(A.fp))
        return out

    def sky(self, sky_desc, direction):
        d = np.ascontiguousarray(direction, np.float32)
        out = np.zeros(3, np.float32)
        self.lib.orc_kat_sky(C.byref(sky_desc), d.ctypes.data_as(A.fp), out.ctypes.data_as(A.fp))
        return out

    def sample_surface(self, desc, material, seed, direction, normal, tangent, uv):
        arrs = [np.ascontiguousarray(x, np.float32) for x in (direction, normal, tangent, uv)]
        wi, thr = np.zeros(3, np.float32), np.zeros(3, np.float32)
        draws = self.lib.orc_kat_sample_surface(C.byref(desc), material, seed, *[a.ctypes.data_as(A.fp) for a in arrs],
                                                wi.ctypes.data_as(A.fp), thr.ctypes.data_as(A.fp))
        return wi, thr, draws

    def intersect_full(self, desc, ray):
        r = np.ascontiguousarray(ray, np.float32).reshape(8)
        out = np.zeros(12, np.float32)
        prim = self.lib.orc_intersect_full(C.byref(desc), r.ctypes.data, out.ctypes.data_as(A.fp))
        return prim, out
