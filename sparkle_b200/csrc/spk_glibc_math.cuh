/*
 * spk_glibc_math.cuh — sinf, cosf and powf as glibc computes them, on the device.
 *
 * The reference calls std::sin / std::cos / std::pow on floats (lens and BxDF sampling, sRGB decode and encode, Schlick's
 * (1 - cos)^5), i.e. glibc's float routines (sysdeps/ieee754/flt-32/s_sinf.c, s_cosf.c, sincosf.h, e_powf.c; the algorithms
 * and tables of ARM's optimized-routines, in glibc since 2.28): a double-precision polynomial after a table-driven
 * reduction, rounded once to float. They are within 0.56 ulp, NOT correctly rounded, so "evaluate in double and round"
 * (what rounds 1 and 2 did) differs from them by one ulp for 1.3 % of the sin / cos arguments and 0.07 % of the pow
 * arguments, and a path that lands on the other side of a branch because of it is a different pixel. Restated here
 * operation for operation, with the fused multiply-adds of the x86-64 FMA build the dynamic loader selects on the CPUs of
 * this pool (sysdeps/x86_64/fpu/multiarch: the same C source compiled with -mfma -mavx2, every a * b + c contracted).
 * tools/glibc_math_probe.cpp checks the restatement on the host against the box's libm: 0 differences in 5e7 arguments
 * each for sin, cos (0..2 pi and +-100), pow(x, 2.4), pow(x, 0.45); without the contractions 24 and 2 in 5e7 differ.
 * tests/test_libm_gpu.py checks the device functions against the same libm, argument by argument.
 * Tables: spk_glibc_tables.h (read out of the image's libm.so.6 by tools/gen_glibc_tables.py).
 * It is also cheaper than CUDA's double sincos() / exp(log()): ~25 double operations per call instead of 100-500.
 */
#pragma once
#include "spk_glibc_tables.h"

namespace spk
{
namespace glibc
{
__device__ __forceinline__ unsigned abstop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ffu; }

// __sincosf_table entry: sign[4], hpi_inv, hpi, c0, c1, s1, c2, s2, c3, s3, c4
// sinf_poly (sincosf.h): sine for even n, cosine for odd n
__device__ __forceinline__ float sinf_poly(double x, double x2, const double *p, int n)
{
    if ((n & 1) == 0)
    {
        const double x3 = x * x2;
        const double s1 = fma(x2, __ldg(p + 12), __ldg(p + 10));
        const double x7 = x3 * x2;
        const double s = fma(x3, __ldg(p + 8), x);
        return (float)fma(x7, s1, s);
    }
    const double x4 = x2 * x2;
    const double c2 = fma(x2, __ldg(p + 13), __ldg(p + 11));
    const double c1 = fma(x2, __ldg(p + 7), __ldg(p + 6));
    const double x6 = x4 * x2;
    const double c = fma(x4, __ldg(p + 9), c1);
    return (float)fma(x6, c2, c);
}

// reduce_fast (sincosf.h): |x| < 120; hpi_inv is 2 / pi * 2^24, so the quadrant ends up in bits 24..31
__device__ __forceinline__ double reduce_fast(double x, int &n)
{
    const double r = x * 0x1.45f306dc9c883p+23;
    n = (__double2int_rz(r) + 0x800000) >> 24;
    return fma(-(double)n, 0x1.921fb54442d18p+0, x);
}

// quadrant -> (is cosine, table); shared by sinf (flip = 0) and cosf (flip = 1)
__device__ __forceinline__ float sincos_one(float y, int flip)
{
    double x = (double)y;
    if (abstop12(y) < abstop12(0x1.921FB6p-1f)) // |y| < pi / 4
    {
        if (abstop12(y) < abstop12(0x1p-12f))
        {
            return flip ? 1.0f : y;
        }
        return sinf_poly(x, x * x, kSinCos0, flip);
    }
    if (abstop12(y) < abstop12(120.0f))
    {
        int n;
        x = reduce_fast(x, n);
        const double s = (n & 3) == 1 || (n & 3) == 2 ? -1.0 : 1.0; // sign[n & 3] = {1, -1, -1, 1}
        const double *p = (n & 2) ? kSinCos1 : kSinCos0;
        return sinf_poly(x * s, x * x, p, n ^ flip);
    }
    // huge arguments (reduce_large) never reach the path tracer: correctly rounded instead
    return (float)(flip ? cos(x) : sin(x));
}
} // namespace glibc

static __device__ __noinline__ float glibc_sinf(float y) { return glibc::sincos_one(y, 0); }
static __device__ __noinline__ float glibc_cosf(float y) { return glibc::sincos_one(y, 1); }
static __device__ __noinline__ void glibc_sincosf(float y, float *s, float *c)
{
    *s = glibc::sincos_one(y, 0);
    *c = glibc::sincos_one(y, 1);
}

// __powf (e_powf.c) for y > 0 finite and x >= 0: log2_inline, y * log2(x), exp2_inline
static __device__ __noinline__ float glibc_powf(float x, float y)
{
    using namespace glibc;
    unsigned ix = __float_as_uint(x);
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) // zero, subnormal, negative, infinite, NaN
    {
        if (ix == 0u)
        {
            return 0.f; // x^y for +0 and y > 0
        }
        if (ix >= 0x00800000u)
        {
            return (float)pow((double)x, (double)y); // negative, infinite, NaN: not on the path tracer's paths
        }
        ix = __float_as_uint(x * 0x1p23f) & 0x7fffffffu; // normalise a subnormal: the exponent becomes negative
        ix -= 23u << 23;
    }
    // log2_inline
    const unsigned tmp = ix - 0x3f330000u;
    const int i = (int)((tmp >> (23 - 4)) % 16u);
    const unsigned top = tmp & 0xff800000u;
    const unsigned iz = ix - top;
    const int k = (int)top >> 23; // arithmetic shift
    const double invc = __ldg(kPowLog2Tab + 2 * i), logc = __ldg(kPowLog2Tab + 2 * i + 1);
    const double z = (double)__uint_as_float(iz);
    const double r = fma(z, invc, -1.0);
    const double y0 = logc + (double)k;
    const double r2 = r * r;
    double q = fma(0x1.27616c9496e0bp-2, r, -0x1.71969a075c67ap-2);  // A[0] r + A[1]
    const double p = fma(0x1.ec70a6ca7baddp-2, r, -0x1.7154748bef6c8p-1); // A[2] r + A[3]
    const double r4 = r2 * r2;
    double t = fma(0x1.71547652ab82bp+0, r, y0); // A[4] r + y0
    t = fma(p, r2, t);
    const double logx = fma(q, r4, t);
    const double ylogx = (double)y * logx;
    if (((unsigned long long)__double_as_longlong(ylogx) >> 47 & 0xffffull) >=
        ((unsigned long long)__double_as_longlong(126.0) >> 47)) // |y log2 x| >= 126: overflow / underflow handling
    {
        return (float)pow((double)x, (double)y);
    }
    // exp2_inline
    double kd = ylogx + kExp2ShiftScaled;
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd -= kExp2ShiftScaled;
    const double rr = ylogx - kd;
    unsigned long long tb = __ldg(kExp2Tab + (ki % 32ull));
    tb += ki << (52 - 5);
    const double s = __longlong_as_double((long long)tb);
    const double zz = fma(0x1.c6af84b912394p-5, rr, 0x1.ebfce50fac4f3p-3); // C[0] r + C[1]
    const double rr2 = rr * rr;
    double out = fma(0x1.62e42ff0c52d6p-1, rr, 1.0); // C[2] r + 1
    out = fma(zz, rr2, out);
    return (float)(out * s);
}
} // namespace spk
