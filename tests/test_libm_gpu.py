"""sin, cos and pow on the device are glibc's float routines restated (csrc/spk_glibc_math.cuh): the reference calls
std::sin / std::cos / std::pow on floats, so the same argument must give the SAME FLOAT as the libm of the box, bit for bit.
Through the C ABI (spk_cuda_probe_libm), against libm.so.6 itself, over the argument ranges the path tracer uses and beyond."""
import ctypes as C
import ctypes.util

import numpy as np
import pytest

from sparkle_b200 import _abi as A
from sparkle_b200.device import CudaContext

pytestmark = pytest.mark.gpu


# The device restates the FMA build of glibc's routines, which the dynamic loader picks on every CPU with FMA (all hosts of
# this pool: 0 differences). On a CPU without FMA libm runs the same source uncontracted and ~5 arguments in 10^7 differ by
# one ulp (tools/glibc_math_probe.cpp -DFMA=0): tolerated, so that such a host does not turn the suite red.
FEW = 3 if "fma" not in open("/proc/cpuinfo").read() else 0


def host_libm():
    m = C.CDLL(ctypes.util.find_library("m") or "libm.so.6")
    for f in (m.sinf, m.cosf):
        f.restype, f.argtypes = C.c_float, [C.c_float]
    m.powf.restype, m.powf.argtypes = C.c_float, [C.c_float, C.c_float]
    return m


def device(ctx, op, x, y=None):
    x = np.ascontiguousarray(x, np.float32)
    out = np.empty_like(x)
    yp = np.ascontiguousarray(y, np.float32).ctypes.data if y is not None else None
    A.check(ctx.lib.spk_cuda_probe_libm(ctx.handle, op, x.ctypes.data, yp, x.size, out.ctypes.data))
    return out


def test_device_sin_cos_pow_equal_glibc():
    g = np.random.default_rng(5)
    n = 400_000
    m = host_libm()
    ctx = CudaContext(16, 16)
    # angles of the samplers (0 .. 2 pi), the whole fast range (|x| < 120), the tiny-argument branches, the quadrant edges
    angles = np.concatenate([g.uniform(0, 2 * np.pi, n), g.uniform(-119.9, 119.9, n), g.uniform(-1e-3, 1e-3, 1000),
                             np.arange(-8, 9) * (np.pi / 4), [0.0, 1e-5, 0.785398, 0.7853982, 2.0 ** -12]]).astype(np.float32)
    for op, f, name in ((0, m.sinf, "sinf"), (1, m.cosf, "cosf")):
        got = device(ctx, op, angles)
        want = np.array([f(float(a)) for a in angles], np.float32)
        bad = np.nonzero(got.view(np.uint32) != want.view(np.uint32))[0]
        assert bad.size <= FEW and np.all(np.abs(got.view(np.int32)[bad] - want.view(np.int32)[bad]) <= 1), \
            (name, bad.size, angles[bad[:5]], got[bad[:5]], want[bad[:5]])
    # sRGB decode ((c + 0.055) / 1.055)^2.4 on (0, 1], encode r^0.45 on [0, 1] incl. 0 and subnormals, and other exponents
    xs = np.concatenate([g.uniform(0.04, 1.0, n), g.uniform(0, 1.0, n), [0.0, 1.0, 1e-39, 1e-45, 0.5, 2.0, 100.0]]).astype(np.float32)
    for yv in (2.4, 0.45, 1.0 / 2.4, 5.0):
        ys = np.full(xs.shape, yv, np.float32)
        got = device(ctx, 2, xs, ys)
        want = np.array([m.powf(float(a), float(np.float32(yv))) for a in xs], np.float32)
        bad = np.nonzero(got.view(np.uint32) != want.view(np.uint32))[0]
        assert bad.size <= FEW and np.all(np.abs(got.view(np.int32)[bad] - want.view(np.int32)[bad]) <= 1), \
            (yv, bad.size, xs[bad[:5]], got[bad[:5]], want[bad[:5]])
    # the special cases of e_powf.c for a positive non-integer exponent: zeros, subnormals, overflow, underflow (down to a
    # subnormal result), infinity, NaN, negative bases (domain error)
    edge = np.array([0.0, -0.0, 1e-45, 1e-40, 1.1754944e-38, 1e-30, 1e-20, 1e-16, 1e16, 1e20, 1e30, 3.4e38, np.inf, np.nan,
                     -1.0, -0.5, -np.inf, 1.0, 0.99999994, 1.0000001], np.float32)
    for yv in (2.4, 0.45):
        got = device(ctx, 2, edge, np.full(edge.shape, yv, np.float32))
        want = np.array([m.powf(float(a), float(np.float32(yv))) for a in edge], np.float32)
        same = (got.view(np.uint32) == want.view(np.uint32)) | (np.isnan(got) & np.isnan(want))
        assert same.all(), (yv, edge[~same], got[~same], want[~same])
    # sin / cos beyond the fast range fall back to a correctly rounded value: within one ulp of glibc's
    big = np.array([120.0, 1000.0, -5000.5, 1e6], np.float32)
    for op, f in ((0, m.sinf), (1, m.cosf)):
        got = device(ctx, op, big)
        want = np.array([f(float(a)) for a in big], np.float32)
        assert np.all(np.abs(got.view(np.int32) - want.view(np.int32)) <= 1), (got, want)
    ctx.close()
