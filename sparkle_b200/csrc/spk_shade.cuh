/*
 * spk_shade.cuh — per-(pixel,sample) RNG, hit attributes, materials, BxDF sampling and the sky, operation for operation
 * as the reference CPU path tracer evaluates them. Compiled with -fmad=false.
 *
 *   RNG            core/math/Sampler.h:16-53, thirdparty/xoshiro_cpp/XoshiroCpp.hpp (Xoshiro128Plus, SplitMix64)
 *   hit attributes MeshRenderProxy::GetIntersection (MeshRenderProxy.cpp:340-361), io/Mesh.h:86-134,
 *                  SphereRenderProxy::GetIntersection (SphereRenderProxy.h:22-31)
 *   textures       Image2D::Sample (io/Image.h:112-152), utilities::SRGBtoLinear (Utilities.h:416-425)
 *   materials      MaterialRenderProxy.h:117-162, PbrMaterialRenderProxy.h:15-51, DialetricMaterialRenderProxy.h:16-38
 *   BxDFs          scene/material/BxDF.h:32-141, DistributionVn (Sampler.h:133-165), Utilities.h:201-324,434-451
 *   sky            SkyRenderProxy::Evaluate (SkyRenderProxy.cpp:50-64), Image2DCube (io/Image.cpp:298-357)
 *
 * libm: sqrt and division are IEEE on the device. sin, cos and pow are glibc's own algorithms (spk_glibc_math.cuh): bit-equal
 * to the libm the reference runs on. (Rounds 1 and 2 evaluated them in double and rounded once, which differs from glibc by
 * one ulp for 1.3 % of the sin / cos and 0.07 % of the pow arguments.)
 */
#pragma once
#include "spk_device.cuh"
#include "spk_glibc_math.cuh"

#include <cuda_fp16.h>

namespace spk
{
constexpr float kEps = 1e-6f;
constexpr float kTolerance = 1e-4f;
constexpr float kPi = 3.14159274101257324f;      // std::numbers::pi_v<float>
constexpr float kInvSqrt3 = 0.577350258827209473f; // std::numbers::inv_sqrt3_v<float>

// sin, cos and pow are glibc's float routines restated on the device (spk_glibc_math.cuh): the same float for the same
// argument as the reference's std::sin / std::cos / std::pow. One out-of-line copy each (three sin/cos pairs and fourteen
// pow call sites would otherwise be inlined into every kernel).
static __device__ __forceinline__ void ref_sincos(float x, float *s, float *c)
{
    const float2 sc = glibc_sincosf(x);
    *s = sc.x;
    *c = sc.y;
}
static __device__ __forceinline__ float ref_pow(float x, float y)
{
    return glibc_powf(x, y);
}
// std::pow(x, 5.f) of Schlick's approximation: the exact product in double, rounded once (glibc's powf agrees except for
// ~0.1 % of the arguments, by one ulp, in a factor no branch depends on; the full routine would cost 10x as much here)
__device__ __forceinline__ float ref_pow5(float x)
{
#ifdef SPK_POW5_GLIBC
    const float r = glibc_powf(fabsf(x), 5.f); // odd integer exponent: the sign of x carries over (e_powf.c sign_bias)
    return signbit(x) ? -r : r;
#else
    double d = (double)x, d2 = d * d;
    return (float)(d2 * d2 * d);
#endif
}

// ---- XoshiroCpp::Xoshiro128Plus seeded through SplitMix64, sampler::RandomUnit ----
struct Rng
{
    uint32_t s0, s1, s2, s3;
};

__device__ __forceinline__ uint32_t splitmix_next(unsigned long long &x)
{
    x += 0x9E3779B97F4A7C15ull;
    unsigned long long z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return (uint32_t)(z ^ (z >> 31));
}

#ifdef SPK_NOINLINE_SEED
static __device__ __noinline__ Rng rng_seed(uint32_t seed)
#else
__device__ __forceinline__ Rng rng_seed(uint32_t seed)
#endif
{
    unsigned long long x = seed;
    Rng r;
    r.s0 = splitmix_next(x);
    r.s1 = splitmix_next(x);
    r.s2 = splitmix_next(x);
    r.s3 = splitmix_next(x);
    return r;
}

__device__ __forceinline__ float rng_unit(Rng &r)
{
    uint32_t result = r.s0 + r.s3;
    uint32_t t = r.s1 << 9;
    r.s2 ^= r.s0;
    r.s3 ^= r.s1;
    r.s1 ^= r.s2;
    r.s0 ^= r.s3;
    r.s2 ^= t;
    r.s3 = (r.s3 << 11) | (r.s3 >> 21);
    // static_cast<float>(r) / static_cast<float>(UINT32_MAX): both round to nearest, the divisor is 2^32, and 1.0f is
    // reachable. Dividing by 2^32 and multiplying by 2^-32 are the same exact exponent shift (no result is subnormal).
    return __uint2float_rn(result) * 2.3283064365386962890625e-10f;
}

// sampler::UnitDisk: rand, then theta = (u * 2.0f) * Pi
__device__ __forceinline__ void unit_disk(Rng &r, float &x, float &y)
{
    float rand = rng_unit(r);
    float theta = rng_unit(r) * 2.0f * kPi;
    float rad = sqrtf(rand);
    float sn, cs;
    ref_sincos(theta, &sn, &cs);
    x = rad * cs;
    y = rad * sn;
}

// ---- textures: Image2D::Sample ----
// utilities::WrapMod (Utilities.h:526-529): ((unsigned)(a % (int)b) + b) % b is the mathematical a mod b for every
// int a (the unsigned wrap cancels); one signed remainder and a conditional add give the same value
__device__ __forceinline__ uint32_t wrap_mod(int a, uint32_t b)
{
    int m = a % (int)b;
    return (uint32_t)(m < 0 ? m + (int)b : m);
}

// ColorToVec = cast<float>() / 255.f (io/Image.h:60-110): the 256 possible IEEE quotients float(k) / 255.f as exact
// hex-float literals (generated with numpy float32 division; tests/test_render_gpu.py checks the texture path against the
// reference), instead of 16 inlined division sequences per bilinear sample
static __device__ const float g_unorm8[256] = {
    0x0.0p+0f, 0x1.0101020000000p-8f, 0x1.0101020000000p-7f, 0x1.8181820000000p-7f, 0x1.0101020000000p-6f, 0x1.4141420000000p-6f,
    0x1.8181820000000p-6f, 0x1.c1c1c20000000p-6f, 0x1.0101020000000p-5f, 0x1.2121220000000p-5f, 0x1.4141420000000p-5f, 0x1.6161620000000p-5f,
    0x1.8181820000000p-5f, 0x1.a1a1a20000000p-5f, 0x1.c1c1c20000000p-5f, 0x1.e1e1e20000000p-5f, 0x1.0101020000000p-4f, 0x1.1111120000000p-4f,
    0x1.2121220000000p-4f, 0x1.3131320000000p-4f, 0x1.4141420000000p-4f, 0x1.5151520000000p-4f, 0x1.6161620000000p-4f, 0x1.7171720000000p-4f,
    0x1.8181820000000p-4f, 0x1.9191920000000p-4f, 0x1.a1a1a20000000p-4f, 0x1.b1b1b20000000p-4f, 0x1.c1c1c20000000p-4f, 0x1.d1d1d20000000p-4f,
    0x1.e1e1e20000000p-4f, 0x1.f1f1f20000000p-4f, 0x1.0101020000000p-3f, 0x1.09090a0000000p-3f, 0x1.1111120000000p-3f, 0x1.19191a0000000p-3f,
    0x1.2121220000000p-3f, 0x1.29292a0000000p-3f, 0x1.3131320000000p-3f, 0x1.39393a0000000p-3f, 0x1.4141420000000p-3f, 0x1.49494a0000000p-3f,
    0x1.5151520000000p-3f, 0x1.59595a0000000p-3f, 0x1.6161620000000p-3f, 0x1.69696a0000000p-3f, 0x1.7171720000000p-3f, 0x1.79797a0000000p-3f,
    0x1.8181820000000p-3f, 0x1.89898a0000000p-3f, 0x1.9191920000000p-3f, 0x1.99999a0000000p-3f, 0x1.a1a1a20000000p-3f, 0x1.a9a9aa0000000p-3f,
    0x1.b1b1b20000000p-3f, 0x1.b9b9ba0000000p-3f, 0x1.c1c1c20000000p-3f, 0x1.c9c9ca0000000p-3f, 0x1.d1d1d20000000p-3f, 0x1.d9d9da0000000p-3f,
    0x1.e1e1e20000000p-3f, 0x1.e9e9ea0000000p-3f, 0x1.f1f1f20000000p-3f, 0x1.f9f9fa0000000p-3f, 0x1.0101020000000p-2f, 0x1.0505060000000p-2f,
    0x1.09090a0000000p-2f, 0x1.0d0d0e0000000p-2f, 0x1.1111120000000p-2f, 0x1.1515160000000p-2f, 0x1.19191a0000000p-2f, 0x1.1d1d1e0000000p-2f,
    0x1.2121220000000p-2f, 0x1.2525260000000p-2f, 0x1.29292a0000000p-2f, 0x1.2d2d2e0000000p-2f, 0x1.3131320000000p-2f, 0x1.3535360000000p-2f,
    0x1.39393a0000000p-2f, 0x1.3d3d3e0000000p-2f, 0x1.4141420000000p-2f, 0x1.4545460000000p-2f, 0x1.49494a0000000p-2f, 0x1.4d4d4e0000000p-2f,
    0x1.5151520000000p-2f, 0x1.5555560000000p-2f, 0x1.59595a0000000p-2f, 0x1.5d5d5e0000000p-2f, 0x1.6161620000000p-2f, 0x1.6565660000000p-2f,
    0x1.69696a0000000p-2f, 0x1.6d6d6e0000000p-2f, 0x1.7171720000000p-2f, 0x1.7575760000000p-2f, 0x1.79797a0000000p-2f, 0x1.7d7d7e0000000p-2f,
    0x1.8181820000000p-2f, 0x1.8585860000000p-2f, 0x1.89898a0000000p-2f, 0x1.8d8d8e0000000p-2f, 0x1.9191920000000p-2f, 0x1.9595960000000p-2f,
    0x1.99999a0000000p-2f, 0x1.9d9d9e0000000p-2f, 0x1.a1a1a20000000p-2f, 0x1.a5a5a60000000p-2f, 0x1.a9a9aa0000000p-2f, 0x1.adadae0000000p-2f,
    0x1.b1b1b20000000p-2f, 0x1.b5b5b60000000p-2f, 0x1.b9b9ba0000000p-2f, 0x1.bdbdbe0000000p-2f, 0x1.c1c1c20000000p-2f, 0x1.c5c5c60000000p-2f,
    0x1.c9c9ca0000000p-2f, 0x1.cdcdce0000000p-2f, 0x1.d1d1d20000000p-2f, 0x1.d5d5d60000000p-2f, 0x1.d9d9da0000000p-2f, 0x1.ddddde0000000p-2f,
    0x1.e1e1e20000000p-2f, 0x1.e5e5e60000000p-2f, 0x1.e9e9ea0000000p-2f, 0x1.ededee0000000p-2f, 0x1.f1f1f20000000p-2f, 0x1.f5f5f60000000p-2f,
    0x1.f9f9fa0000000p-2f, 0x1.fdfdfe0000000p-2f, 0x1.0101020000000p-1f, 0x1.0303040000000p-1f, 0x1.0505060000000p-1f, 0x1.0707080000000p-1f,
    0x1.09090a0000000p-1f, 0x1.0b0b0c0000000p-1f, 0x1.0d0d0e0000000p-1f, 0x1.0f0f100000000p-1f, 0x1.1111120000000p-1f, 0x1.1313140000000p-1f,
    0x1.1515160000000p-1f, 0x1.1717180000000p-1f, 0x1.19191a0000000p-1f, 0x1.1b1b1c0000000p-1f, 0x1.1d1d1e0000000p-1f, 0x1.1f1f200000000p-1f,
    0x1.2121220000000p-1f, 0x1.2323240000000p-1f, 0x1.2525260000000p-1f, 0x1.2727280000000p-1f, 0x1.29292a0000000p-1f, 0x1.2b2b2c0000000p-1f,
    0x1.2d2d2e0000000p-1f, 0x1.2f2f300000000p-1f, 0x1.3131320000000p-1f, 0x1.3333340000000p-1f, 0x1.3535360000000p-1f, 0x1.3737380000000p-1f,
    0x1.39393a0000000p-1f, 0x1.3b3b3c0000000p-1f, 0x1.3d3d3e0000000p-1f, 0x1.3f3f400000000p-1f, 0x1.4141420000000p-1f, 0x1.4343440000000p-1f,
    0x1.4545460000000p-1f, 0x1.4747480000000p-1f, 0x1.49494a0000000p-1f, 0x1.4b4b4c0000000p-1f, 0x1.4d4d4e0000000p-1f, 0x1.4f4f500000000p-1f,
    0x1.5151520000000p-1f, 0x1.5353540000000p-1f, 0x1.5555560000000p-1f, 0x1.5757580000000p-1f, 0x1.59595a0000000p-1f, 0x1.5b5b5c0000000p-1f,
    0x1.5d5d5e0000000p-1f, 0x1.5f5f600000000p-1f, 0x1.6161620000000p-1f, 0x1.6363640000000p-1f, 0x1.6565660000000p-1f, 0x1.6767680000000p-1f,
    0x1.69696a0000000p-1f, 0x1.6b6b6c0000000p-1f, 0x1.6d6d6e0000000p-1f, 0x1.6f6f700000000p-1f, 0x1.7171720000000p-1f, 0x1.7373740000000p-1f,
    0x1.7575760000000p-1f, 0x1.7777780000000p-1f, 0x1.79797a0000000p-1f, 0x1.7b7b7c0000000p-1f, 0x1.7d7d7e0000000p-1f, 0x1.7f7f800000000p-1f,
    0x1.8181820000000p-1f, 0x1.8383840000000p-1f, 0x1.8585860000000p-1f, 0x1.8787880000000p-1f, 0x1.89898a0000000p-1f, 0x1.8b8b8c0000000p-1f,
    0x1.8d8d8e0000000p-1f, 0x1.8f8f900000000p-1f, 0x1.9191920000000p-1f, 0x1.9393940000000p-1f, 0x1.9595960000000p-1f, 0x1.9797980000000p-1f,
    0x1.99999a0000000p-1f, 0x1.9b9b9c0000000p-1f, 0x1.9d9d9e0000000p-1f, 0x1.9f9fa00000000p-1f, 0x1.a1a1a20000000p-1f, 0x1.a3a3a40000000p-1f,
    0x1.a5a5a60000000p-1f, 0x1.a7a7a80000000p-1f, 0x1.a9a9aa0000000p-1f, 0x1.ababac0000000p-1f, 0x1.adadae0000000p-1f, 0x1.afafb00000000p-1f,
    0x1.b1b1b20000000p-1f, 0x1.b3b3b40000000p-1f, 0x1.b5b5b60000000p-1f, 0x1.b7b7b80000000p-1f, 0x1.b9b9ba0000000p-1f, 0x1.bbbbbc0000000p-1f,
    0x1.bdbdbe0000000p-1f, 0x1.bfbfc00000000p-1f, 0x1.c1c1c20000000p-1f, 0x1.c3c3c40000000p-1f, 0x1.c5c5c60000000p-1f, 0x1.c7c7c80000000p-1f,
    0x1.c9c9ca0000000p-1f, 0x1.cbcbcc0000000p-1f, 0x1.cdcdce0000000p-1f, 0x1.cfcfd00000000p-1f, 0x1.d1d1d20000000p-1f, 0x1.d3d3d40000000p-1f,
    0x1.d5d5d60000000p-1f, 0x1.d7d7d80000000p-1f, 0x1.d9d9da0000000p-1f, 0x1.dbdbdc0000000p-1f, 0x1.ddddde0000000p-1f, 0x1.dfdfe00000000p-1f,
    0x1.e1e1e20000000p-1f, 0x1.e3e3e40000000p-1f, 0x1.e5e5e60000000p-1f, 0x1.e7e7e80000000p-1f, 0x1.e9e9ea0000000p-1f, 0x1.ebebec0000000p-1f,
    0x1.ededee0000000p-1f, 0x1.efeff00000000p-1f, 0x1.f1f1f20000000p-1f, 0x1.f3f3f40000000p-1f, 0x1.f5f5f60000000p-1f, 0x1.f7f7f80000000p-1f,
    0x1.f9f9fa0000000p-1f, 0x1.fbfbfc0000000p-1f, 0x1.fdfdfe0000000p-1f, 0x1.0000000000000p+0f,
};

__device__ __forceinline__ float4 access_pixel(const DTexture &t, uint32_t x, uint32_t y)
{
    size_t idx = size_t(y) * t.width + x;
    if (t.format <= 1u) // RGBA8 unorm / srgb
    {
        uchar4 c = __ldg(reinterpret_cast<const uchar4 *>(t.pixels) + idx);
        return make_float4(g_unorm8[c.x], g_unorm8[c.y], g_unorm8[c.z], g_unorm8[c.w]);
    }
    if (t.format == 2u) // RGBAFloat16
    {
        uint2 raw = __ldg(reinterpret_cast<const uint2 *>(t.pixels) + idx);
        __half2 a = *reinterpret_cast<__half2 *>(&raw.x), b = *reinterpret_cast<__half2 *>(&raw.y);
        float2 fa = __half22float2(a), fb = __half22float2(b);
        return make_float4(fa.x, fa.y, fb.x, fb.y);
    }
    return __ldg(reinterpret_cast<const float4 *>(t.pixels) + idx);
}

__device__ __forceinline__ float lerp1(float a, float b, float f) { return a + (b - a) * f; }

__device__ __forceinline__ float srgb_to_linear(float c)
{
    float cutoff = c < 0.04045f ? 1.f : 0.f;
    float higher = ref_pow((c + 0.055f) / 1.055f, 2.4f);
    float lower = c / 12.92f;
    return higher * (1.f - cutoff) + lower * cutoff;
}

static __device__ __noinline__ V3 sample_texture(const DTexture &t, float u, float v)
{
    float px = u * (float)(t.width - 1), py = v * (float)(t.height - 1);
    int ui = (int)px, vi = (int)py; // utilities::Decompose: truncation toward zero
    float uc = px - (float)ui, vc = py - (float)vi;
    // WrapMod(ui + 1) = (WrapMod(ui) + 1) mod width: a compare instead of a second remainder
    uint32_t x0 = wrap_mod(ui, t.width), x1 = x0 + 1u == t.width ? 0u : x0 + 1u;
    uint32_t y0 = wrap_mod(vi, t.height), y1 = y0 + 1u == t.height ? 0u : y0 + 1u;
    float4 s00 = access_pixel(t, x0, y0), s10 = access_pixel(t, x1, y0);
    float4 s01 = access_pixel(t, x0, y1), s11 = access_pixel(t, x1, y1);
    float r0 = lerp1(s00.x, s10.x, uc), g0 = lerp1(s00.y, s10.y, uc), b0 = lerp1(s00.z, s10.z, uc);
    float r1 = lerp1(s01.x, s11.x, uc), g1 = lerp1(s01.y, s11.y, uc), b1 = lerp1(s01.z, s11.z, uc);
    V3 c = mk(lerp1(r0, r1, vc), lerp1(g0, g1, vc), lerp1(b0, b1, vc));
    if (t.format == 1u) // pseudo sRGB decode after filtering
    {
        // (three powers in one out-of-line call, for the scheduler to interleave, measured no faster: 31.9 vs 31.5 ms)
        c = mk(srgb_to_linear(c.x), srgb_to_linear(c.y), srgb_to_linear(c.z));
    }
    return c;
}

// utilities::GetPossibleMajorAxis
__device__ __forceinline__ V3 major_axis(V3 n)
{
    if (fabsf(n.x) < kInvSqrt3)
    {
        return mk(1.f, 0.f, 0.f);
    }
    if (fabsf(n.y) < kInvSqrt3)
    {
        return mk(0.f, 1.f, 0.f);
    }
    return mk(0.f, 0.f, 1.f);
}

// ---- hit attributes: {Mesh,Sphere}RenderProxy::GetIntersection ----
struct Surface
{
    V3 location, normal, tangent;
    float u, v; // texture coordinate
    uint32_t material;
};

template <typename T> __device__ __forceinline__ float bary(T v0, T v1, T v2, float u, float v)
{
    return v0 + (v1 - v0) * u + (v2 - v0) * v; // utilities::Lerp(v0, v1, v2, u, v)
}

#ifdef SPK_NOINLINE_NMAP
static __device__ __noinline__ void material_normal_texture(const DScene &sc, const DMaterial &m, float tu, float tv,
                                                            V3 tangent, float headedness, V3 &n)
#else
__device__ __forceinline__ void material_normal_texture(const DScene &sc, const DMaterial &m, float tu, float tv,
                                                        V3 tangent, float headedness, V3 &n)
#endif
{
    // MaterialRenderProxy::GetNormal = Sample(uv) * 2 - Ones; utilities::TangentSpaceToWorldSpace
    V3 s = sample_texture(sc.textures[m.normal_tex], tu, tv);
    V3 tn = mk(s.x * 2.f - 1.f, s.y * 2.f - 1.f, s.z * 2.f - 1.f);
    // NB `auto bi_tangent = ...` (Utilities.h:403-409) is an Eigen expression over a temporary: undefined behaviour in
    // the reference. This follows what the source says; the oracle TU is built so that it does too (oracle/Makefile).
    V3 bi = cross(n, tangent) * headedness;
    n = normalized(mk(tangent.x * tn.x + bi.x * tn.y + n.x * tn.z, tangent.y * tn.x + bi.y * tn.y + n.y * tn.z,
                      tangent.z * tn.x + bi.z * tn.y + n.z * tn.z));
}

__device__ __forceinline__ void get_intersection(const DScene &sc, V3 wo, V3 wd, const Hit &hit, Surface &sf)
{
    const DInstance *__restrict__ in = sc.instances + hit.inst;
    sf.material = __ldg(&in->material);
    sf.location = wo + wd * hit.t; // Intersection::Update -> Ray::At(t)
    if (__ldg(&in->kind) == 1u)
    {
        V3 center = mk(__ldg(&in->fwd[3]), __ldg(&in->fwd[7]), __ldg(&in->fwd[11]));
        V3 p = wo + wd * hit.t;
        sf.normal = normalized(p - center);
        sf.tangent = major_axis(sf.normal);
        sf.u = 0.f;
        sf.v = 0.f;
        return;
    }
    const DMesh &mesh = sc.meshes[__ldg(&in->mesh)];
    const uint32_t *idx = mesh.indices + size_t(hit.face) * 3;
    uint32_t i0 = __ldg(idx), i1 = __ldg(idx + 1), i2 = __ldg(idx + 2);
    float u = hit.u, v = hit.v;
    if (mesh.uvs)
    {
        const float2 *t = reinterpret_cast<const float2 *>(mesh.uvs);
        float2 t0 = __ldg(t + i0), t1 = __ldg(t + i1), t2 = __ldg(t + i2);
        sf.u = bary(t0.x, t1.x, t2.x, u, v);
        sf.v = bary(t0.y, t1.y, t2.y, u, v);
    }
    else
    {
        sf.u = 0.f;
        sf.v = 0.f;
    }
    const float *nn = mesh.normals;
    V3 n = mk(bary(__ldg(nn + 3 * i0), __ldg(nn + 3 * i1), __ldg(nn + 3 * i2), u, v),
              bary(__ldg(nn + 3 * i0 + 1), __ldg(nn + 3 * i1 + 1), __ldg(nn + 3 * i2 + 1), u, v),
              bary(__ldg(nn + 3 * i0 + 2), __ldg(nn + 3 * i1 + 2), __ldg(nn + 3 * i2 + 2), u, v));
    const float4 *tt = reinterpret_cast<const float4 *>(mesh.tangents);
    float4 g0 = __ldg(tt + i0), g1 = __ldg(tt + i1), g2 = __ldg(tt + i2);
    V3 tg = mk(bary(g0.x, g1.x, g2.x, u, v), bary(g0.y, g1.y, g2.y, u, v), bary(g0.z, g1.z, g2.z, u, v));
    float tw = bary(g0.w, g1.w, g2.w, u, v);
    const DMaterial &m = sc.materials[sf.material];
    if (__ldg(&m.normal_tex) != SPK_INVALID_ID)
    {
        material_normal_texture(sc, m, sf.u, sf.v, tg, tw, n);
    }
    // inv_transform.TransformDirectionTangentSpace = the FORWARD 3x3 block (Transform.h:105-108 on the inverse)
    float fwd[12];
    const float4 *mm = reinterpret_cast<const float4 *>(in) + 3;
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        float4 a = __ldg(mm + k);
        fwd[4 * k] = a.x;
        fwd[4 * k + 1] = a.y;
        fwd[4 * k + 2] = a.z;
        fwd[4 * k + 3] = a.w;
    }
    sf.normal = normalized(xform_dir(fwd, n));
    sf.tangent = xform_dir(fwd, tg);
}

// ---- MaterialRenderProxy getters ----
struct MaterialSample
{
    V3 base_color;
    float roughness, metallic;
};

__device__ __forceinline__ V3 material_emissive(const DScene &sc, const DMaterial &m, float u, float v)
{
    V3 e = mk(__ldg(&m.emissive[0]), __ldg(&m.emissive[1]), __ldg(&m.emissive[2]));
    uint32_t tex = __ldg(&m.emissive_tex);
    if (tex != SPK_INVALID_ID)
    {
        e = cmul(sample_texture(sc.textures[tex], u, v), e);
    }
    return e;
}

__device__ __forceinline__ MaterialSample material_sample(const DScene &sc, const DMaterial &m, float u, float v)
{
    MaterialSample s;
    s.base_color = mk(__ldg(&m.base_color[0]), __ldg(&m.base_color[1]), __ldg(&m.base_color[2]));
    s.roughness = __ldg(&m.roughness);
    s.metallic = __ldg(&m.metallic);
    uint32_t bt = __ldg(&m.base_color_tex), mt = __ldg(&m.metallic_roughness_tex);
    if (bt != SPK_INVALID_ID)
    {
        s.base_color = cmul(sample_texture(sc.textures[bt], u, v), s.base_color);
    }
    if (mt != SPK_INVALID_ID)
    {
        V3 mr = sample_texture(sc.textures[mt], u, v);
        s.roughness = mr.y * s.roughness;
        s.metallic = mr.z * s.metallic;
    }
    return s;
}

// ---- BxDF helpers (Utilities.h) ----
__device__ __forceinline__ float schlick1(float cos_i, float r0) { return r0 + (1.f - r0) * ref_pow5(1.f - cos_i); }

__device__ __forceinline__ float smith_ggx_correlated(float cos_o, float cos_i, float roughness)
{
    float a = roughness * roughness, a2 = a * a;
    float ggx_i = cos_o * sqrtf((-cos_i * a2 + cos_i) * cos_i + a2);
    float ggx_o = cos_i * sqrtf((-cos_o * a2 + cos_o) * cos_o + a2);
    return 2.f * cos_o * cos_i / (ggx_o + ggx_i);
}

__device__ __forceinline__ float geometry_schlick_ggx(float cos_theta, float roughness)
{
    float a = roughness * roughness, a2 = a * a;
    float tan2 = (1.f - cos_theta * cos_theta) / (cos_theta * cos_theta);
    return 2.f / (1.f + sqrtf(1.f + a2 * tan2));
}

// sampler::DistributionVn::Sample (Sampler.h:133-165)
__device__ __forceinline__ V3 sample_vndf(Rng &rng, V3 wo, float roughness)
{
    float a = roughness * roughness;
    float u1 = rng_unit(rng), u2 = rng_unit(rng);
    V3 v = normalized(mk(wo.x * a, wo.y * a, wo.z));
    V3 t1 = v.z < 0.999f ? normalized(cross(v, mk(0.f, 0.f, 1.f))) : mk(1.f, 0.f, 0.f);
    V3 t2 = cross(v, t1);
    float r = sqrtf(u1);
    float phi = 2.f * kPi * u2;
    float sn, cs;
    ref_sincos(phi, &sn, &cs);
    float p1 = r * cs, p2 = r * sn;
    float s = 0.5f * (1.f + v.z);
    p2 = (1.f - s) * sqrtf(1.f - p1 * p1) + s * p2;
    float c = sqrtf(fmaxf(0.f, 1.f - p1 * p1 - p2 * p2));
    V3 n = mk(p1 * t1.x + p2 * t2.x + c * v.x, p1 * t1.y + p2 * t2.y + c * v.y, p1 * t1.z + p2 * t2.z + c * v.z);
    return normalized(mk(a * n.x, a * n.y, fmaxf(0.f, n.z)));
}

// utilities::GetLocalAxisFromNormal(normal, major_axis = tangent)
__device__ __forceinline__ void local_axes(V3 normal, V3 tangent, V3 &u, V3 &v, V3 &w)
{
    u = normalized(cross(normal, tangent));
    v = cross(normal, u);
    w = normal;
}

// MaterialRenderProxy::SampleSurface for Pbr / Dielectric; returns this bounce's throughput, sets wi (world)
// DieletricBxDF::Sample (BxDF.h:112-141), FrDielectric (Utilities.h:301-324), EtaI = 1
#ifdef SPK_NOINLINE_DIEL
static __device__ __noinline__ V3 dielectric_sample(float eta, V3 lwo, Rng &rng)
#else
__device__ __forceinline__ V3 dielectric_sample(float eta, V3 lwo, Rng &rng)
#endif
{
    V3 lwi;
    float cos_i = lwo.z, eta_i = 1.f, eta_t = eta;
    if (!(cos_i > 0.f))
    {
        eta_i = eta;
        eta_t = 1.f;
        cos_i = -cos_i;
    }
    float sin_i = sqrtf(1.f - cos_i * cos_i);
    float ref_idx = eta_i / eta_t;
    float sin_t = ref_idx * sin_i;
    float fresnel;
    if (sin_t >= 1.f)
    {
        fresnel = 1.f;
    }
    else
    {
        float r0 = (1.f - ref_idx) / (1.f + ref_idx);
        r0 = r0 * r0;
        fresnel = schlick1(cos_i, r0);
    }
    if (rng_unit(rng) < fresnel)
    {
        lwi = mk(-lwo.x, -lwo.y, lwo.z); // utilities::Reflect(w)
    }
    else
    {
        bool entering = lwo.z > 0.f;
        float ratio = entering ? 1.f / eta : eta / 1.f;
        // utilities::Refract
        float ct = lwo.z;
        float sin_i2 = 1.f - ct * ct;
        float sin_t2 = ratio * ratio * sin_i2;
        float cos_t = sqrtf(1.f - sin_t2) * (ct > 0.f ? -1.f : 1.f);
        lwi = normalized(mk(-ratio * lwo.x, -ratio * lwo.y, cos_t));
    }
    return lwi;
}

__device__ __forceinline__ V3 sample_surface(const DScene &sc, Rng &rng, const Surface &sf, V3 ray_dir, V3 &wi_world)
{
    const DMaterial &m = sc.materials[sf.material];
    MaterialSample ms = material_sample(sc, m, sf.u, sf.v);
    V3 au, av, aw;
    local_axes(sf.normal, sf.tangent, au, av, aw);
    V3 w_o = -ray_dir;
    V3 lwo = mk(dot(w_o, au), dot(w_o, av), dot(w_o, aw)); // TransformBasisToLocal
    V3 lwi = mk(0.f, 0.f, 0.f);
    V3 throughput;
    if (__ldg(&m.type) == 1u)
    {
        lwi = dielectric_sample(__ldg(&m.eta), lwo, rng);
        throughput = ms.base_color;
    }
    else
    {
        // SpecularBxDF::Sample (BxDF.h:47-110)
        V3 wm = sample_vndf(rng, lwo, ms.roughness);
        float k = dot(lwo, wm) * 2.f;
        V3 swi = mk(wm.x * k - lwo.x, wm.y * k - lwo.y, wm.z * k - lwo.z); // utilities::Reflect(w_o, w_m)
        float cos_o = fmaxf(fabsf(lwo.z), 0.f), cos_in = fmaxf(fabsf(swi.z), 0.f), cos_m = fmaxf(fabsf(wm.z), 0.f);
        float local_cos_i = dot(swi, wm);
        bool valid = false;
        V3 fresnel_color = mk(0.f, 0.f, 0.f);
        if (cos_o > kEps && cos_in > kEps && cos_m > kEps)
        {
            if (rng_unit(rng) < ms.metallic)
            {
                float f = ref_pow5(1.f - local_cos_i);
                fresnel_color = mk(ms.base_color.x + (1.f - ms.base_color.x) * f,
                                   ms.base_color.y + (1.f - ms.base_color.y) * f,
                                   ms.base_color.z + (1.f - ms.base_color.z) * f);
                valid = true;
            }
            else
            {
                float fresnel = schlick1(local_cos_i, 0.04f);
                if (rng_unit(rng) < fresnel)
                {
                    fresnel_color = mk(1.f, 1.f, 1.f);
                    valid = true;
                }
            }
        }
        if (valid)
        {
            float occlusion = smith_ggx_correlated(cos_o, cos_in, ms.roughness);
            float normalizer = geometry_schlick_ggx(cos_o, ms.roughness) + kEps;
            throughput = mk(fresnel_color.x * occlusion / normalizer, fresnel_color.y * occlusion / normalizer,
                            fresnel_color.z * occlusion / normalizer);
            lwi = swi;
        }
        else
        {
            // LambertianBxDF::Sample -> CosineWeightedHemiSphere::Sample
            float dx, dy;
            unit_disk(rng, dx, dy);
            float z = sqrtf(fmaxf(0.f, 1.0f - (dx * dx + dy * dy)));
            lwi = mk(dx, dy, z);
            throughput = ms.base_color;
        }
    }
    // TransformBasisToWorld(...).normalized()
    wi_world = normalized(mk(au.x * lwi.x + av.x * lwi.y + aw.x * lwi.z, au.y * lwi.x + av.y * lwi.y + aw.y * lwi.z,
                             au.z * lwi.x + av.z * lwi.y + aw.z * lwi.z));
    return throughput;
}

// material class of a traced segment: the key warps are regrouped by before shading (wf_shade, render_sorted_kernel)
__device__ __forceinline__ uint32_t shade_class(const DScene &sc, uint32_t inst)
{
    // 0 miss; else 1 + mesh + 2 dielectric + 4 textured + 8 emissive, precomputed per instance at upload (spk_api.cu)
    return inst == SPK_INVALID_ID ? 0u : __ldg(&sc.instances[inst].shade_class); // 0..16
}

// ---- SkyRenderProxy::Evaluate ----
__device__ __forceinline__ V3 sky_evaluate(const DScene &sc, V3 d)
{
    if (sc.sky.has_map)
    {
        // Image2DCube::DirectionToTextureCoordinate (io/Image.cpp:298-345); FaceId order +X,-X,+Y,-Y,+Z,-Z
        float ax = fabsf(d.x), ay = fabsf(d.y), az = fabsf(d.z), u, v;
        int face;
        if (ax >= ay && ax >= az)
        {
            if (d.x > 0.f) { face = 0; u = -d.z / d.x; v = -d.y / d.x; }
            else { face = 1; u = d.z / -d.x; v = -d.y / -d.x; }
        }
        else if (ay >= ax && ay >= az)
        {
            if (d.y > 0.f) { face = 2; u = d.x / d.y; v = d.z / d.y; }
            else { face = 3; u = d.x / -d.y; v = -d.z / -d.y; }
        }
        else
        {
            if (d.z > 0.f) { face = 4; u = d.x / d.z; v = -d.y / d.z; }
            else { face = 5; u = -d.x / -d.z; v = -d.y / -d.z; }
        }
        u = (u + 1.f) * 0.5f;
        v = (v + 1.f) * 0.5f;
        return sample_texture(sc.textures[sc.sky.face_base + face], u, v);
    }
    float t = 0.5f * d.z + 0.5f;
    return mk(1.f + (sc.sky.color[0] - 1.f) * t, 1.f + (sc.sky.color[1] - 1.f) * t, 1.f + (sc.sky.color[2] - 1.f) * t);
}
} // namespace spk
