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
 * It is also cheaper than CUDA's double sincos() / exp(log()) once written branch-free with immediate coefficients and
 * without calls into CUDA's double routines (measured steps in DESIGN.md 3): TestScene 32.73 -> 32.08 ms per 64 spp.
 */
#pragma once
#include "spk_glibc_tables.h"

namespace spk
{
namespace glibc
{
__device__ __forceinline__ unsigned abstop12(float x) { return (__float_as_uint(x) >> 20) & 0x7ffu; }

// sinf / cosf / sincosf for |y| < 120 (s_sinf.c, s_cosf.c, sincosf.h), branch-free: lanes of a warp fall into different
// quadrants and argument ranges, and a literal transcription runs the sine and the cosine polynomial for each of them in
// turn. What the three routines compute is the same two polynomials:
//   reduce_fast: n = round(y * 2/pi) (through the 2^24-scaled product and an arithmetic shift), x = y - n * pi/2; for
//                |y| < pi/4 it gives n = 0 and x = y exactly, which is the routines' separate small-argument branch
//   sinf_poly:   S = x s + (x s)^3 s1 + (x s)^5 (s2 + x^2 s3)   with s = sign[n & 3] = {1, -1, -1, 1}
//                C = (c0 + x^2 c1) + x^4 c2 + x^6 (c3 + x^2 c4)
//   __sincosf_table[1] (quadrants 2, 3) is table [0] with the cosine coefficients negated: every term of C is linear in them
//   and rounding is symmetric, so C's value is exactly negated
//   sin y = n odd ? C : S, cos y = n odd ? S : C; below 2^-12 sin y = y and cos y = 1
// Coefficients are immediates (a table in global memory put a trip to L1 in front of every polynomial: TestScene -25 %).
__device__ __forceinline__ void sincos_fast(float y, float &sin_out, float &cos_out)
{
    const double x0 = (double)y;
    const double r = x0 * 0x1.45f306dc9c883p+23; // hpi_inv = 2 / pi * 2^24
    const int n = (__double2int_rz(r) + 0x800000) >> 24;
    const double x = fma(-(double)n, 0x1.921fb54442d18p+0, x0);
    const double xs = ((n + 1) & 2) ? -x : x;
    const double x2 = x * x;
    // sine polynomial
    const double x3 = xs * x2;
    const double s1 = fma(x2, -0x1.994eb3774cf24p-13, 0x1.1107605230bc4p-7); // s2 + x2 s3
    const double x5 = x3 * x2;
    const double s = fma(x3, -0x1.555545995a603p-3, xs); // x + x3 s1
    const float S = (float)fma(x5, s1, s);
    // cosine polynomial
    const double x4 = x2 * x2;
    const double c2 = fma(x2, 0x1.99343027bf8c3p-16, -0x1.6c087e89a359dp-10); // c3 + x2 c4
    const double c1 = fma(x2, -0x1.ffffffd0c621cp-2, 1.0);                    // c0 + x2 c1
    const double x6 = x4 * x2;
    const double c = fma(x4, 0x1.55553e1068f19p-5, c1); // c1 + x4 c2
    float C = (float)fma(x6, c2, c);
    C = (n & 2) ? -C : C;
    const bool tiny = abstop12(y) < abstop12(0x1p-12f);
    sin_out = tiny ? y : ((n & 1) ? C : S);
    cos_out = tiny ? 1.0f : ((n & 1) ? S : C);
}
} // namespace glibc

// |y| >= 120 (glibc: reduce_large) never reaches the path tracer: correctly rounded instead. A function of its own: CUDA's
// double sin / cos carry a Payne-Hanek path whose registers the fast routine would otherwise be charged with.
static __device__ __noinline__ float2 glibc_sincosf_huge(float y)
{
    return make_float2((float)sin((double)y), (float)cos((double)y));
}
// {sin y, cos y}, returned in registers: with pointer arguments the caller's two floats lived in local memory (a store and a
// load through L1 per call, since the function is not inlined)
#ifdef SPK_INLINE_SINCOS
static __device__ __forceinline__ float2 glibc_sincosf(float y)
#else
static __device__ __noinline__ float2 glibc_sincosf(float y)
#endif
{
    if (glibc::abstop12(y) < glibc::abstop12(120.0f))
    {
        float2 r;
        glibc::sincos_fast(y, r.x, r.y);
        return r;
    }
    return glibc_sincosf_huge(y);
}

// __powf (e_powf.c) for finite y > 0 that is not an integer (2.4, 0.45, 1 / 2.4: the sRGB curves): log2_inline,
// y * log2(x), exp2_inline, with the routine's own handling of zero, subnormal, infinite and NaN x and of overflow and
// underflow; a negative x is a domain error for such y (NaN). No call into CUDA's double pow(): its registers would be
// charged to every caller.
namespace glibc
{
__device__ __forceinline__ float powf_core(float x, float y)
{
    unsigned ix = __float_as_uint(x);
    if (ix - 0x00800000u >= 0x7f800000u - 0x00800000u) // zero, subnormal, negative, infinite, NaN
    {
        if (ix == 0u)
        {
            return 0.f; // x^y for +0 and y > 0
        }
        if (ix >= 0x7f800000u)
        {
            // +-inf -> x * x = +inf; -0 -> +0 (y is not an odd integer); finite x < 0: domain error; NaN stays NaN
            return (ix & 0x7fffffffu) == 0x7f800000u ? __int_as_float(0x7f800000) : (ix == 0x80000000u ? 0.f : __int_as_float(0x7fc00000));
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
    const double invc = SPK_GLIBC_LD(kPowLog2Tab + 2 * i), logc = SPK_GLIBC_LD(kPowLog2Tab + 2 * i + 1);
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
        if (ylogx > 0x1.fffffffd1d571p+6)
        {
            return __int_as_float(0x7f800000); // overflow
        }
        if (ylogx <= -150.0)
        {
            return 0.f; // underflow
        }
        // between: a subnormal result, rounded by the conversion at the end like glibc's
    }
    // exp2_inline
    double kd = ylogx + kExp2ShiftScaled;
    const unsigned long long ki = (unsigned long long)__double_as_longlong(kd);
    kd -= kExp2ShiftScaled;
    const double rr = ylogx - kd;
    unsigned long long tb = SPK_GLIBC_LD(kExp2Tab + (ki % 32ull));
    tb += ki << (52 - 5);
    const double s = __longlong_as_double((long long)tb);
    const double zz = fma(0x1.c6af84b912394p-5, rr, 0x1.ebfce50fac4f3p-3); // C[0] r + C[1]
    const double rr2 = rr * rr;
    double out = fma(0x1.62e42ff0c52d6p-1, rr, 1.0); // C[2] r + 1
    out = fma(zz, rr2, out);
    return (float)(out * s);
}
} // namespace glibc

static __device__ __noinline__ float glibc_powf(float x, float y)
{
    return glibc::powf_core(x, y);
}
} // namespace spk
