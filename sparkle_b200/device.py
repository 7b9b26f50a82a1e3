"""Thin Python handle over the C ABI (include/spk_cuda.h) for tests, bench.py and the headless driver.
Every call goes through libsparkle_cuda.so; there is no fallback of any kind."""
import ctypes as C

import numpy as np

from . import _abi as A

RAY_DTYPE = np.dtype([("origin", np.float32, 3), ("pad0", np.float32), ("direction", np.float32, 3),
                      ("pad1", np.float32)])
HIT_DTYPE = np.dtype([("t", np.float32), ("u", np.float32), ("v", np.float32), ("prim", np.uint32),
                      ("face", np.uint32)])
assert RAY_DTYPE.itemsize == C.sizeof(A.Ray) and HIT_DTYPE.itemsize == C.sizeof(A.Hit)


class CudaContext:
    def __init__(self, width, height, device=0):
        self.lib = A.cuda_lib()
        self.width, self.height = width, height
        info = A.CreateInfo(A.ABI_VERSION, device, width, height)
        self.handle = A.CtxP()
        A.check(self.lib.spk_cuda_create(C.byref(info), C.byref(self.handle)))

    def close(self):
        if self.handle:
            self.lib.spk_cuda_destroy(self.handle)
            self.handle = A.CtxP()
            for p in getattr(self, "_pinned", []):
                self.lib.spk_cuda_host_free(p)
            self._pinned = []

    def __del__(self):
        try:
            self.close()
        except Exception:
            pass

    def set_stream(self, cuda_stream):
        A.check(self.lib.spk_cuda_set_stream(self.handle, C.c_void_p(cuda_stream)))

    def set_option(self, option, value):
        """SpkOption tuning knobs (A.OPT_*): explicit, nothing is read from the environment"""
        A.check(self.lib.spk_cuda_set_option(self.handle, option, int(value)))

    def upload_scene(self, desc):
        A.check(self.lib.spk_cuda_upload_scene(self.handle, C.byref(desc)))

    def update_instances(self, prim_array, n, tlas_desc):
        A.check(self.lib.spk_cuda_update_instances(self.handle, prim_array, n, C.byref(tlas_desc)))

    def set_camera(self, cam):
        A.check(self.lib.spk_cuda_set_camera(self.handle, C.byref(cam)))

    def reset_accumulation(self):
        A.check(self.lib.spk_cuda_reset_accumulation(self.handle))

    def render(self, first_sample, sample_count, max_bounce=8, sample_stride=1, output_limit=6.0,
               accum_mode=A.ACCUM_SUM, flags=0, cumulated_cap=0):
        p = A.RenderParams(first_sample, sample_count, sample_stride, max_bounce, output_limit, accum_mode, flags,
                           cumulated_cap)
        A.check(self.lib.spk_cuda_render(self.handle, C.byref(p)))

    RESOLVE_DTYPES = {A.RESOLVE_LINEAR_F32: np.float32, A.RESOLVE_TONEMAP_F16: np.float16,
                      A.RESOLVE_TONEMAP_SRGB8: np.uint8}

    def resolve(self, fmt=A.RESOLVE_LINEAR_F32, total_samples=0, out=None):
        """tone-map + read back; `out` may be a pinned frame buffer from pinned_frame() (one DMA, no staging)"""
        if out is None:
            out = np.empty((self.height, self.width, 4), self.RESOLVE_DTYPES[fmt])
        A.check(self.lib.spk_cuda_resolve(self.handle, fmt, total_samples, out.ctypes.data, out.nbytes))
        return out

    def pinned_frame(self, fmt=A.RESOLVE_TONEMAP_F16):
        """(H, W, 4) array over page-locked host memory (spk_cuda_host_alloc); freed with the context"""
        dt = np.dtype(self.RESOLVE_DTYPES[fmt])
        nbytes = self.height * self.width * 4 * dt.itemsize
        p = C.c_void_p()
        A.check(self.lib.spk_cuda_host_alloc(C.byref(p), nbytes))
        self._pinned = getattr(self, "_pinned", []) + [p]
        buf = (C.c_uint8 * nbytes).from_address(p.value)
        return np.frombuffer(buf, dtype=dt).reshape(self.height, self.width, 4)

    def accum_ptr(self):
        p, n = C.c_void_p(), C.c_size_t()
        A.check(self.lib.spk_cuda_accum_device_ptr(self.handle, C.byref(p), C.byref(n)))
        return p.value, n.value

    def accum_sample_count(self):
        n = A.u32()
        A.check(self.lib.spk_cuda_accum_sample_count(self.handle, C.byref(n)))
        return n.value

    def trace_closest(self, rays, flags=0):
        """rays: (n, 8) float32 {o.xyz, pad, d.xyz, pad} in HOST memory; H2D + traversal + D2H inside the call"""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        hits = np.empty(len(rays), HIT_DTYPE)
        A.check(self.lib.spk_cuda_trace_closest(self.handle, rays.ctypes.data, len(rays), hits.ctypes.data, flags))
        return hits

    def trace_closest_noalloc(self, rays, flags=0, hits=None):
        """trace_closest through the library's static device scratch: no allocation inside the call; at most
        A.NOALLOC_MAX_RAYS rays. `hits`: optional preallocated HIT_DTYPE array to fill."""
        rays = np.ascontiguousarray(rays, np.float32).reshape(-1, 8)
        if hits is None:
            hits = np.empty(len(rays), HIT_DTYPE)
        A.check(self.lib.spk_cuda_trace_closest_noalloc(self.handle, rays.ctypes.data, len(rays), hits.ctypes.data, flags))
        return hits

    def trace_closest_device(self, d_rays, n, d_hits, flags=0):
        A.check(self.lib.spk_cuda_trace_closest_device(self.handle, C.c_void_p(d_rays), n, C.c_void_p(d_hits), flags))

    def choose_tracer(self, max_bounce=8):
        """measure megakernel vs wavefront on the current scene; returns (wavefront?, (ms_mega, ms_wave))"""
        w, ms = C.c_uint32(0), (C.c_float * 2)()
        A.check(self.lib.spk_cuda_choose_tracer(self.handle, max_bounce, C.byref(w), ms))
        return bool(w.value), (ms[0], ms[1])

    def tracer_choice(self):
        w, form = C.c_int32(-1), C.c_uint32(0)
        A.check(self.lib.spk_cuda_tracer_choice(self.handle, C.byref(w), C.byref(form)))
        return w.value, form.value

    def stats(self):
        s = A.Stats()
        A.check(self.lib.spk_cuda_get_stats(self.handle, C.byref(s)))
        return {k: getattr(s, k) for k, _ in A.Stats._fields_}

    def reset_stats(self):
        A.check(self.lib.spk_cuda_reset_stats(self.handle))

    def synchronize(self):
        A.check(self.lib.spk_cuda_synchronize(self.handle))
