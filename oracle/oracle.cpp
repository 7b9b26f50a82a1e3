/*
 * oracle.cpp — TEST INFRASTRUCTURE ONLY (builds oracle/liboracle.so, orc_kind() == "port").
 *
 * Plain C++ restatement of the reference's CPU path tracer, written against the SpkSceneDesc data contract of
 * include/spk_cuda.h. No Eigen, no reference headers, no product code: every reduction spells out the association the
 * reference's Eigen expressions compile to with g++ -O2 -ffp-contract=off, and every function cites the
 * reference file:line it follows (paths relative to /root/reference/libraries unless they start with thirdparty/).
 *
 * PINNING. This port is pinned two ways (tests/test_oracle_cpu.py):
 *   (1) against oracle/_ref/libsparkle_ref.so — the reference's OWN translation units (CPURenderer.cpp,
 *       SceneRenderProxy.cpp, MeshRenderProxy.cpp, BxDF.h, Sampler.h ...) compiled in place — wherever /root/reference is
 *       mounted: RNG streams, Eigen association KATs, texture/sky/BxDF samples, closest-hit records and whole images
 *       must agree BIT FOR BIT on the same inputs;
 *   (2) against tests/golden/*.npz, vectors generated from (1) by tests/golden/make_golden.py and committed, so the
 *       pin travels to machines without the reference.
 * What stays unpinned by any reference-owned test: the bvh::v2 traversal contract (madmann91/bvh @5a9d759c is absent
 * from the mount and the reference has no test at that boundary). Closest-hit results are checked against
 * brute force (topology independent) instead.
 *
 * Build: g++ -std=c++20 -O2 -ffp-contract=off (baseline x86-64: no FMA contraction, SSE2 float math).
 * Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline / --impl reference legs may load this library.
 */
#include <algorithm>
#include <atomic>
#include <cfloat>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

#include "oracle.h"

namespace
{
// ---------------------------------------------------------------------------------------------------------------------
// constants: core/math/Types.h:45-57
constexpr float kEps = 1e-6f;
constexpr float kTolerance = 1e-4f;
constexpr float kPi = 3.14159274101257324f;        // std::numbers::pi_v<float>

// libm mode (orc_set_libm, port only). 0 = glibc's float routines, what the reference calls (default; every parity pin
// against the reference uses it). 1 = "device libm": what the CUDA path computes. Since the device restates glibc's sinf /
// cosf / powf (csrc/spk_glibc_math.cuh) the only difference left is Schlick's (1 - cos)^5, which the device evaluates as an
// exact product in double rounded once (spk_shade.cuh ref_pow5) where the reference calls std::pow(x, 5.f): one ulp apart
// for ~0.1 % of the arguments. Mode 1 lets a test separate "the GPU follows the reference's operation order" (GPU == port
// in mode 1, to accumulation rounding) from that share (port mode 1 vs reference, CPU only). (Until late in round 2 the
// device evaluated sin, cos and pow in double and rounded once, and mode 1 did the same.)
int g_libm_mode = 0;
inline float m_sin(float x) { return std::sin(x); }
inline float m_cos(float x) { return std::cos(x); }
inline float m_pow(float x, float y)
{
    if (g_libm_mode && y == 5.f)
    {
        const double d = static_cast<double>(x), d2 = d * d;
        return static_cast<float>(d2 * d2 * d);
    }
    return std::pow(x, y);
}
constexpr float kInvSqrt3 = 0.577350258827209473f; // std::numbers::inv_sqrt3_v<float>

// ---------------------------------------------------------------------------------------------------------------------
// vector vocabulary with the association of the reference's Eigen expressions (probed by orc_kat_vec)
struct V3
{
    float x, y, z;
};
inline V3 mk(float x, float y, float z) { return V3{x, y, z}; }
inline V3 mk(const float *p) { return V3{p[0], p[1], p[2]}; }
inline V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
inline V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
inline V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
inline V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
inline V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
inline V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
inline V3 cmul(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
// Vector3::dot: (a0*b0 + a1*b1) + a2*b2
inline float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
// Vector3::squaredNorm: a0*a0 + (a1*a1 + a2*a2)
inline float sqnorm(V3 a) { return a.x * a.x + (a.y * a.y + a.z * a.z); }
inline float norm(V3 a) { return std::sqrt(sqnorm(a)); }
// Vector3::normalized: true division by sqrt(squaredNorm)
inline V3 normalized(V3 a)
{
    float n = std::sqrt(sqnorm(a));
    return mk(a.x / n, a.y / n, a.z / n);
}
inline V3 cross(V3 a, V3 b) { return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x); }
// Transform::TransformDirection (core/math/Transform.h:100-103): linear block row = m0*v0 + (m1*v1 + m2*v2)
inline V3 xform_dir(const float *m, V3 v)
{
    return mk(m[0] * v.x + (m[1] * v.y + m[2] * v.z), m[4] * v.x + (m[5] * v.y + m[6] * v.z),
              m[8] * v.x + (m[9] * v.y + m[10] * v.z));
}
// Transform::TransformPoint (core/math/Transform.h:95-98): affine row = ((m0*p0 + m1*p1) + m2*p2) + t
inline V3 xform_point(const float *m, V3 p)
{
    return mk(((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3], ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7],
              ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11]);
}
inline float lerp1(float a, float b, float f) { return a + (b - a) * f; } // utilities::Lerp (Utilities.h:119-122)

// ---------------------------------------------------------------------------------------------------------------------
// a1. RNG: core/math/Sampler.h:16-41; thirdparty/xoshiro_cpp/XoshiroCpp.hpp:684-690 (SplitMix64), :1305-1330
// (Xoshiro128Plus: seeded with four SplitMix64 outputs truncated to 32 bits)
struct Rng
{
    uint32_t s[4];
    uint64_t draws = 0;
};

inline uint32_t splitmix_next(uint64_t &x)
{
    x += 0x9E3779B97F4A7C15ull;
    uint64_t z = x;
    z = (z ^ (z >> 30)) * 0xBF58476D1CE4E5B9ull;
    z = (z ^ (z >> 27)) * 0x94D049BB133111EBull;
    return static_cast<uint32_t>(z ^ (z >> 31));
}

inline Rng rng_seed(uint32_t seed) // sampler::ReseedCurrentThread(seed): the u32 is zero-extended to the u64 seed
{
    uint64_t x = seed;
    Rng r;
    for (int k = 0; k < 4; k++)
    {
        r.s[k] = splitmix_next(x);
    }
    return r;
}

inline uint32_t rng_next(Rng &r)
{
    const uint32_t result = r.s[0] + r.s[3];
    const uint32_t t = r.s[1] << 9;
    r.s[2] ^= r.s[0];
    r.s[3] ^= r.s[1];
    r.s[1] ^= r.s[2];
    r.s[0] ^= r.s[3];
    r.s[2] ^= t;
    r.s[3] = (r.s[3] << 11) | (r.s[3] >> 21);
    r.draws++;
    return result;
}

// sampler::RandomUnit (Sampler.h:32-41): float(r) / float(UINT32_MAX); both conversions round to nearest, so the
// divisor is 2^32 and 1.0f is reachable
inline float rng_unit(Rng &r) { return static_cast<float>(rng_next(r)) / static_cast<float>(UINT32_MAX); }

// sampler::UnitDisk (Sampler.h:43-53): rand first, then theta = u * 2.0f * Pi
inline void unit_disk(Rng &r, float &x, float &y)
{
    float rand = rng_unit(r);
    float theta = rng_unit(r) * 2.0f * kPi;
    float rad = std::sqrt(rand);
    x = rad * m_cos(theta);
    y = rad * m_sin(theta);
}

// ---------------------------------------------------------------------------------------------------------------------
// a11. textures: Image2D::Sample (io/Image.h:112-152), AccessPixel/ColorToVec (io/Image.h:60-110),
// utilities::WrapMod (Utilities.h:526-529), utilities::SRGBtoLinear (Utilities.h:416-425)
inline uint32_t wrap_mod(int a, uint32_t b) { return (static_cast<uint32_t>(a % static_cast<int>(b)) + b) % b; }

inline float half_to_float(uint16_t h) // IEEE binary16 -> binary32 (Eigen::half conversion is exact)
{
    uint32_t sign = (h >> 15) & 1u, exp = (h >> 10) & 31u, man = h & 1023u, bits;
    if (exp == 0)
    {
        if (man == 0)
        {
            bits = sign << 31;
        }
        else
        {
            int e = -1;
            do
            {
                e++;
                man <<= 1;
            } while ((man & 1024u) == 0);
            bits = (sign << 31) | (uint32_t(127 - 15 - e) << 23) | ((man & 1023u) << 13);
        }
    }
    else if (exp == 31)
    {
        bits = (sign << 31) | 0x7F800000u | (man << 13);
    }
    else
    {
        bits = (sign << 31) | ((exp + 112u) << 23) | (man << 13);
    }
    float f;
    memcpy(&f, &bits, 4);
    return f;
}

inline uint16_t float_to_half(float f) // round to nearest even, as Eigen::half(float)
{
    uint32_t x;
    memcpy(&x, &f, 4);
    uint32_t sign = (x >> 16) & 0x8000u;
    x &= 0x7FFFFFFFu;
    if (x >= 0x7F800000u)
    {
        return uint16_t(sign | (x > 0x7F800000u ? 0x7E00u : 0x7C00u));
    }
    if (x >= 0x477FF000u) // rounds to >= 65520: infinity
    {
        return uint16_t(sign | 0x7C00u);
    }
    if (x < 0x38800000u) // subnormal half or zero
    {
        if (x < 0x33000000u)
        {
            return uint16_t(sign);
        }
        uint32_t e = x >> 23, m = (x & 0x7FFFFFu) | 0x800000u;
        uint32_t shift = 126u - e; // 14..24
        uint32_t r = m >> shift, rem = m & ((1u << shift) - 1u), halfway = 1u << (shift - 1);
        if (rem > halfway || (rem == halfway && (r & 1u)))
        {
            r++;
        }
        return uint16_t(sign | r);
    }
    uint32_t r = ((x - 0x38000000u) >> 13), rem = x & 0x1FFFu;
    if (rem > 0x1000u || (rem == 0x1000u && (r & 1u)))
    {
        r++;
    }
    return uint16_t(sign | r);
}

struct F4
{
    float x, y, z, w;
};

inline F4 access_pixel(const SpkTextureDesc &t, uint32_t x, uint32_t y)
{
    size_t idx = size_t(y) * t.width + x;
    if (t.format == SPK_FMT_RGBA8_UNORM || t.format == SPK_FMT_RGBA8_SRGB) // ColorToVec: cast<float>() / 255.f
    {
        const uint8_t *c = static_cast<const uint8_t *>(t.pixels) + idx * 4;
        return F4{float(c[0]) / 255.f, float(c[1]) / 255.f, float(c[2]) / 255.f, float(c[3]) / 255.f};
    }
    if (t.format == SPK_FMT_RGBA16F)
    {
        const uint16_t *c = static_cast<const uint16_t *>(t.pixels) + idx * 4;
        return F4{half_to_float(c[0]), half_to_float(c[1]), half_to_float(c[2]), half_to_float(c[3])};
    }
    const float *c = static_cast<const float *>(t.pixels) + idx * 4;
    return F4{c[0], c[1], c[2], c[3]};
}

inline float srgb_to_linear(float c)
{
    float cutoff = c < 0.04045f ? 1.f : 0.f;
    float higher = m_pow((c + 0.055f) / 1.055f, 2.4f);
    float lower = c / 12.92f;
    return higher * (1.f - cutoff) + lower * cutoff;
}

V3 sample_texture(const SpkTextureDesc &t, float u, float v)
{
    float px = u * float(t.width - 1), py = v * float(t.height - 1);
    int ui = int(px), vi = int(py); // utilities::Decompose: C cast, truncation toward zero; the fraction may be negative
    float uc = px - float(ui), vc = py - float(vi);
    uint32_t x0 = wrap_mod(ui, t.width), x1 = wrap_mod(ui + 1, t.width);
    uint32_t y0 = wrap_mod(vi, t.height), y1 = wrap_mod(vi + 1, t.height);
    F4 s00 = access_pixel(t, x0, y0), s10 = access_pixel(t, x1, y0);
    F4 s01 = access_pixel(t, x0, y1), s11 = access_pixel(t, x1, y1);
    float r0 = lerp1(s00.x, s10.x, uc), g0 = lerp1(s00.y, s10.y, uc), b0 = lerp1(s00.z, s10.z, uc);
    float r1 = lerp1(s01.x, s11.x, uc), g1 = lerp1(s01.y, s11.y, uc), b1 = lerp1(s01.z, s11.z, uc);
    V3 c = mk(lerp1(r0, r1, vc), lerp1(g0, g1, vc), lerp1(b0, b1, vc));
    if (t.format == SPK_FMT_RGBA8_SRGB) // decode AFTER filtering (io/Image.h:146-150)
    {
        c = mk(srgb_to_linear(c.x), srgb_to_linear(c.y), srgb_to_linear(c.z));
    }
    return c;
}

// ---------------------------------------------------------------------------------------------------------------------
// Scene view over the borrowed SpkSceneDesc. Triangle records are precomputed once per mesh exactly as
// BLAS::Build / Triangle::Set do (renderer/proxy/MeshRenderProxy.cpp:31-65, 78-92): p0 = v0, e1 = v1 - v0, e2 = v2 - v0.
struct Tree
{
    const uint8_t *nodes = nullptr;
    uint32_t num_nodes = 0, stride = 32;
    const void *prim_ids = nullptr;
    uint32_t id_bytes = 4, num_prims = 0;

    void From(const SpkBvhDesc &d)
    {
        nodes = static_cast<const uint8_t *>(d.nodes);
        num_nodes = d.num_nodes;
        stride = d.node_stride;
        prim_ids = d.prim_ids;
        id_bytes = d.prim_id_bytes;
        num_prims = d.num_prims;
    }
    const float *Bounds(uint32_t n) const { return reinterpret_cast<const float *>(nodes + size_t(n) * stride); }
    uint32_t Index(uint32_t n) const
    {
        uint32_t v;
        memcpy(&v, nodes + size_t(n) * stride + 24, 4);
        return v;
    }
    uint32_t PrimId(uint32_t slot) const
    {
        return id_bytes == 8 ? uint32_t(static_cast<const uint64_t *>(prim_ids)[slot])
                             : static_cast<const uint32_t *>(prim_ids)[slot];
    }
};

struct TriRecord
{
    V3 p0, e1, e2;
};

struct MeshView
{
    const SpkMeshDesc *desc = nullptr;
    Tree blas;
    std::vector<TriRecord> tris; // face order
};

struct SceneView
{
    const SpkSceneDesc *desc = nullptr;
    Tree tlas;
    std::vector<MeshView> meshes;
};

struct Counters
{
    uint64_t rays = 0, node_visits = 0, tri_tests = 0, inst_visits = 0, sphere_tests = 0, rng_draws = 0;
    void AddTo(OrcCounters &o) const
    {
        o.rays += rays;
        o.node_visits += node_visits;
        o.tri_tests += tri_tests;
        o.inst_visits += inst_visits;
        o.sphere_tests += sphere_tests;
        o.rng_draws += rng_draws;
    }
};

void build_view(const SpkSceneDesc *d, SceneView &sv)
{
    sv.desc = d;
    sv.tlas.From(d->tlas);
    sv.meshes.resize(d->num_meshes);
    for (uint32_t m = 0; m < d->num_meshes; m++)
    {
        const SpkMeshDesc &md = d->meshes[m];
        MeshView &mv = sv.meshes[m];
        mv.desc = &md;
        mv.blas.From(md.blas);
        mv.tris.resize(md.num_faces);
        for (uint32_t f = 0; f < md.num_faces; f++)
        {
            V3 v0 = mk(md.vertices + 3 * size_t(md.indices[3 * size_t(f)]));
            V3 v1 = mk(md.vertices + 3 * size_t(md.indices[3 * size_t(f) + 1]));
            V3 v2 = mk(md.vertices + 3 * size_t(md.indices[3 * size_t(f) + 2]));
            mv.tris[f] = TriRecord{v0, v1 - v0, v2 - v0};
        }
    }
}

// IntersectionCandidate (core/math/Intersection.h:91-104): t starts at FLT_MAX
struct Candidate
{
    float t = std::numeric_limits<float>::max();
    float u = 0.f, v = 0.f;
    uint32_t face = SPK_INVALID_ID;
    uint32_t prim = SPK_INVALID_ID; // SceneRenderProxy registration order
};

// ---------------------------------------------------------------------------------------------------------------------
// bvh::v2::Bvh::intersect<IsAnyHit=false, UseRobustTraversal=false> — madmann91/bvh pinned at 5a9d759c (absent from the
// mount: .SUBMODULES.json:65-71; call sites core/math/BVH.h:12-24, SceneRenderProxy.cpp:95-113, MeshRenderProxy.cpp:
// 108-123). Published algorithm restated: inverse direction through safe_inverse (1/x, or copysign(FLT_MAX, x) when
// |x| <= FLT_EPSILON); slab test as a separate multiply and add per bound, near/far picked by the ray octant;
// robust_max/min(a, b) = a > b ? a : b; both children hit -> descend the nearer by entry distance (swap only when
// left.t0 > right.t0), push the farther; leaf -> leaf_fn(first_id, first_id + prim_count). PARITY UNPINNED (no reference
// test at this boundary). CULLED mode additionally skips a child whose entry distance exceeds the current best t.
struct SlabRay
{
    float inv[3], org[3];
    int oct[3];
};

inline float safe_inverse(float x)
{
    return std::fabs(x) <= std::numeric_limits<float>::epsilon() ? std::copysign(std::numeric_limits<float>::max(), x)
                                                                 : 1.f / x;
}

inline SlabRay make_slab_ray(V3 o, V3 d)
{
    SlabRay r;
    const float oo[3] = {o.x, o.y, o.z}, dd[3] = {d.x, d.y, d.z};
    for (int a = 0; a < 3; a++)
    {
        r.inv[a] = safe_inverse(dd[a]);
        r.org[a] = -oo[a] * r.inv[a];
        r.oct[a] = std::signbit(dd[a]) ? 1 : 0;
    }
    return r;
}

inline bool slab(const SlabRay &r, const float *b, float &t0)
{
    float lo = 0.f, hi = std::numeric_limits<float>::max(); // bvh_ray.tmin = 0, tmax = FLT_MAX, never shortened (H2)
    for (int a = 0; a < 3; a++)
    {
        float tn = b[2 * a + r.oct[a]] * r.inv[a] + r.org[a];
        float tf = b[2 * a + 1 - r.oct[a]] * r.inv[a] + r.org[a];
        lo = tn > lo ? tn : lo;
        hi = tf < hi ? tf : hi;
    }
    t0 = lo;
    return lo <= hi;
}

template <typename LeafFn>
void traverse(const Tree &tree, const SlabRay &sr, bool culled, const float &best_t, Counters &cnt, LeafFn &&leaf)
{
    if (tree.num_nodes == 0)
    {
        return;
    }
    uint32_t stack[64]; // SmallStack<Index, 32> in the reference; depth never approaches it on the test scenes
    int sp = 0;
    stack[sp++] = tree.Index(0); // bvh_.get_root().index
    while (sp > 0)
    {
        uint32_t top = stack[--sp];
        bool dead = false;
        while ((top & 15u) == 0u)
        {
            uint32_t first = top >> 4;
            cnt.node_visits++;
            float tl, tr;
            bool hl = slab(sr, tree.Bounds(first), tl), hr = slab(sr, tree.Bounds(first + 1), tr);
            if (culled)
            {
                hl = hl && tl <= best_t;
                hr = hr && tr <= best_t;
            }
            uint32_t li = tree.Index(first), ri = tree.Index(first + 1);
            if (hl)
            {
                uint32_t near_i = li;
                if (hr)
                {
                    uint32_t far_i = ri;
                    if (tl > tr)
                    {
                        std::swap(near_i, far_i);
                    }
                    stack[sp++] = far_i;
                }
                top = near_i;
            }
            else if (hr)
            {
                top = ri;
            }
            else
            {
                dead = true;
                break;
            }
        }
        if (!dead)
        {
            leaf(top >> 4, (top >> 4) + (top & 15u));
        }
    }
}

// ---------------------------------------------------------------------------------------------------------------------
// a8. BLAS::IntersectTriangle<false> (renderer/proxy/MeshRenderProxy.cpp:128-194): Moeller-Trumbore in object space
// (not watertight), then world_t re-derived from the x component of the transformed hit point (Ray::InverseAt,
// core/math/Ray.h:43-47), accepted iff world_t > 0 and world_t < candidate.t (strict).
inline bool intersect_triangle(const TriRecord &tri, uint32_t face, V3 o, V3 d, V3 wo, V3 wd, const float *fwd,
                               Candidate &cand)
{
    V3 h = cross(d, tri.e2);
    float a = dot(tri.e1, h);
    if (std::abs(a) < kEps)
    {
        return false;
    }
    V3 s = o - tri.p0;
    float f = 1.f / a;
    float u = f * dot(s, h);
    if (u < 0.f || u > 1.f)
    {
        return false;
    }
    V3 q = cross(s, tri.e1);
    float v = f * dot(d, q);
    if (v < 0.f || u + v > 1.f)
    {
        return false;
    }
    float t = f * dot(tri.e2, q);
    if (t > kEps)
    {
        V3 lp = o + d * t;              // Ray::At
        V3 wp = xform_point(fwd, lp);   // transform.TransformPoint
        float world_t = (wp.x - wo.x) / wd.x; // Ray::InverseAt: x component only
        if (world_t > 0.f && world_t < cand.t)
        {
            cand.t = world_t;
            cand.u = u;
            cand.v = v;
            cand.face = face;
            return true;
        }
    }
    return false;
}

// a9. SphereRenderProxy::IntersectInternal<false> (renderer/proxy/SphereRenderProxy.cpp:5-38): a = 1 assumed,
// root -sqrt when the origin is outside (c >= 0) else +sqrt; accept t > 0 and t < candidate.t
inline bool intersect_sphere(const SpkPrimitiveDesc &p, V3 wo, V3 wd, Candidate &cand)
{
    V3 center = mk(p.fwd[3], p.fwd[7], p.fwd[11]);
    V3 oc = wo - center;
    float half_b = dot(wd, oc);
    float c = sqnorm(oc) - p.radius2;
    float disc = half_b * half_b - 1.f * c;
    if (disc < 0.f)
    {
        return false;
    }
    float sq = std::sqrt(disc);
    float t = (-half_b + (c >= 0.f ? -sq : sq)) / 1.f;
    if (t > 0.f && t < cand.t)
    {
        cand.t = t;
        return true;
    }
    return false;
}

// a7. BLAS::IntersectInternal<false> (MeshRenderProxy.cpp:94-126): the world ray is moved into object space by the
// inverse transform WITHOUT renormalising the direction (Ray::TransformedBy, core/math/Ray.h:23-26)
bool intersect_mesh(const SceneView &sv, const SpkPrimitiveDesc &p, V3 wo, V3 wd, uint32_t mode, Candidate &cand,
                    Counters &cnt)
{
    const MeshView &mv = sv.meshes[p.mesh];
    cnt.inst_visits++;
    V3 lo = xform_point(p.inv, wo);
    V3 ld = xform_dir(p.inv, wd);
    bool any = false;
    if (mode == ORC_TRACE_BRUTE_FORCE)
    {
        for (uint32_t f = 0; f < mv.desc->num_faces; f++)
        {
            cnt.tri_tests++;
            any |= intersect_triangle(mv.tris[f], f, lo, ld, wo, wd, p.fwd, cand);
        }
        return any;
    }
    SlabRay sr = make_slab_ray(lo, ld);
    traverse(mv.blas, sr, mode == ORC_TRACE_BVH_CULLED, cand.t, cnt, [&](uint32_t begin, uint32_t end) {
        for (uint32_t i = begin; i < end; i++)
        {
            uint32_t face = mv.blas.PrimId(i);
            cnt.tri_tests++;
            any |= intersect_triangle(mv.tris[face], face, lo, ld, wo, wd, p.fwd, cand);
        }
    });
    return any;
}

inline bool intersect_primitive(const SceneView &sv, uint32_t prim, V3 wo, V3 wd, uint32_t mode, Candidate &cand,
                                Counters &cnt)
{
    const SpkPrimitiveDesc &p = sv.desc->primitives[prim];
    bool hit;
    if (p.kind == SPK_PRIM_SPHERE)
    {
        cnt.sphere_tests++;
        hit = intersect_sphere(p, wo, wd, cand);
        if (hit)
        {
            // the sphere test leaves the candidate's u, v, face_idx untouched; GetIntersection ignores them
            // (SphereRenderProxy.h:22-31). Reported as 0 / INVALID like the harness does.
            cand.u = 0.f;
            cand.v = 0.f;
            cand.face = SPK_INVALID_ID;
        }
    }
    else
    {
        hit = intersect_mesh(sv, p, wo, wd, mode, cand, cnt);
    }
    if (hit)
    {
        cand.prim = prim; // candidate.primitive = primitives_[i] (SceneRenderProxy.cpp:103-104)
    }
    return hit;
}

// a6. TLAS::IntersectInternal<false> (renderer/proxy/SceneRenderProxy.cpp:82-126): leaf slot i holds
// primitives_[prim_ids[i]] after the reorder of TLAS::Build (:59-68)
void trace_closest(const SceneView &sv, V3 wo, V3 wd, uint32_t mode, Candidate &cand, Counters &cnt)
{
    cnt.rays++;
    const SpkSceneDesc *d = sv.desc;
    if (mode == ORC_TRACE_BRUTE_FORCE)
    {
        for (uint32_t p = 0; p < d->num_primitives; p++)
        {
            intersect_primitive(sv, p, wo, wd, mode, cand, cnt);
        }
        return;
    }
    if (d->num_primitives == 0)
    {
        return;
    }
    SlabRay sr = make_slab_ray(wo, wd);
    traverse(sv.tlas, sr, mode == ORC_TRACE_BVH_CULLED, cand.t, cnt, [&](uint32_t begin, uint32_t end) {
        for (uint32_t i = begin; i < end; i++)
        {
            intersect_primitive(sv, sv.tlas.PrimId(i), wo, wd, mode, cand, cnt);
        }
    });
}

// ---------------------------------------------------------------------------------------------------------------------
// a10. hit attributes: MeshRenderProxy::GetIntersection (renderer/proxy/MeshRenderProxy.cpp:340-361),
// Mesh::{GetTexCoord,GetShadingNormal,GetTangent} (io/Mesh.h:86-134), SphereRenderProxy::GetIntersection
// (renderer/proxy/SphereRenderProxy.h:22-31), utilities::GetPossibleMajorAxis (Utilities.h:222-234)
struct Surface
{
    V3 location, normal, tangent;
    float u = 0.f, v = 0.f; // texture coordinate
    uint32_t material = 0;
};

inline V3 major_axis(V3 n)
{
    if (std::abs(n.x) < kInvSqrt3)
    {
        return mk(1.f, 0.f, 0.f); // Right
    }
    if (std::abs(n.y) < kInvSqrt3)
    {
        return mk(0.f, 1.f, 0.f); // Front
    }
    return mk(0.f, 0.f, 1.f); // Up
}

// utilities::Lerp(v0, v1, v2, u, v) = v0 + (v1 - v0) * u + (v2 - v0) * v, left to right
inline float bary(float v0, float v1, float v2, float u, float v) { return v0 + (v1 - v0) * u + (v2 - v0) * v; }

void get_intersection(const SceneView &sv, V3 wo, V3 wd, const Candidate &cand, Surface &sf)
{
    const SpkSceneDesc *d = sv.desc;
    const SpkPrimitiveDesc &p = d->primitives[cand.prim];
    sf.material = p.material;
    sf.location = wo + wd * cand.t; // Intersection::Update -> Ray::At(t) with the WORLD ray (Intersection.h:53-57)
    if (p.kind == SPK_PRIM_SPHERE)
    {
        V3 center = mk(p.fwd[3], p.fwd[7], p.fwd[11]);
        sf.normal = normalized(sf.location - center);
        sf.tangent = major_axis(sf.normal);
        sf.u = 0.f;
        sf.v = 0.f;
        return;
    }
    const SpkMeshDesc &m = d->meshes[p.mesh];
    const uint32_t *idx = m.indices + size_t(cand.face) * 3;
    uint32_t i0 = idx[0], i1 = idx[1], i2 = idx[2];
    float u = cand.u, v = cand.v;
    if (m.uvs) // Mesh::GetTexCoord returns Zero when the mesh has no uvs
    {
        sf.u = bary(m.uvs[2 * i0], m.uvs[2 * i1], m.uvs[2 * i2], u, v);
        sf.v = bary(m.uvs[2 * i0 + 1], m.uvs[2 * i1 + 1], m.uvs[2 * i2 + 1], u, v);
    }
    const float *nn = m.normals;
    V3 n = mk(bary(nn[3 * i0], nn[3 * i1], nn[3 * i2], u, v), bary(nn[3 * i0 + 1], nn[3 * i1 + 1], nn[3 * i2 + 1], u, v),
              bary(nn[3 * i0 + 2], nn[3 * i1 + 2], nn[3 * i2 + 2], u, v));
    const float *tt = m.tangents;
    V3 tg = mk(bary(tt[4 * i0], tt[4 * i1], tt[4 * i2], u, v), bary(tt[4 * i0 + 1], tt[4 * i1 + 1], tt[4 * i2 + 1], u, v),
               bary(tt[4 * i0 + 2], tt[4 * i1 + 2], tt[4 * i2 + 2], u, v));
    float tw = bary(tt[4 * i0 + 3], tt[4 * i1 + 3], tt[4 * i2 + 3], u, v);
    const SpkMaterialDesc &mat = d->materials[p.material];
    if (mat.normal_tex != SPK_INVALID_ID)
    {
        // MaterialRenderProxy::GetNormal = Sample(uv) * 2 - Ones (MaterialRenderProxy.h:150-162);
        // utilities::TangentSpaceToWorldSpace (Utilities.h:403-409). NB the reference binds `auto bi_tangent` to an Eigen
        // expression over a temporary (undefined behaviour); this follows what the source says, and the reference TU is
        // compiled so that it does too (oracle/Makefile, MeshRenderProxy.o).
        V3 s = sample_texture(d->textures[mat.normal_tex], sf.u, sf.v);
        V3 tn = mk(s.x * 2.f - 1.f, s.y * 2.f - 1.f, s.z * 2.f - 1.f);
        V3 bi = cross(n, tg) * tw;
        // TransformBasisToWorld (Utilities.h:246-250): u * x + v * y + w * z, left to right
        n = normalized(mk(tg.x * tn.x + bi.x * tn.y + n.x * tn.z, tg.y * tn.x + bi.y * tn.y + n.y * tn.z,
                          tg.z * tn.x + bi.z * tn.y + n.z * tn.z));
    }
    // inv_transform.TransformDirectionTangentSpace uses the inverse's inverse = the FORWARD linear block
    // (Transform.h:105-108), not the inverse transpose
    sf.normal = normalized(xform_dir(p.fwd, n));
    sf.tangent = xform_dir(p.fwd, tg);
}

// ---------------------------------------------------------------------------------------------------------------------
// a11. MaterialRenderProxy::{GetBaseColor,GetRoughness,GetMetallic,GetEmissive} (MaterialRenderProxy.h:117-148):
// factor * texel when the texture exists; roughness = texel.y, metallic = texel.z
inline V3 material_emissive(const SpkSceneDesc *d, const SpkMaterialDesc &m, float u, float v)
{
    V3 e = mk(m.emissive);
    if (m.emissive_tex != SPK_INVALID_ID)
    {
        e = cmul(sample_texture(d->textures[m.emissive_tex], u, v), e);
    }
    return e;
}

struct MaterialSample
{
    V3 base_color;
    float roughness, metallic;
};

inline MaterialSample material_sample(const SpkSceneDesc *d, const SpkMaterialDesc &m, float u, float v)
{
    MaterialSample s{mk(m.base_color), m.roughness, m.metallic};
    if (m.base_color_tex != SPK_INVALID_ID)
    {
        s.base_color = cmul(sample_texture(d->textures[m.base_color_tex], u, v), s.base_color);
    }
    if (m.metallic_roughness_tex != SPK_INVALID_ID)
    {
        V3 mr = sample_texture(d->textures[m.metallic_roughness_tex], u, v);
        s.roughness = mr.y * s.roughness;
        s.metallic = mr.z * s.metallic;
    }
    return s;
}

// ---------------------------------------------------------------------------------------------------------------------
// a12/a13. BxDF helpers: Utilities.h:291-299 (SchlickApproximation), :301-324 (FrDielectric), :434-451 (Smith/Schlick GGX)
inline float schlick(float cos_i, float r0) { return r0 + (1.f - r0) * m_pow(1.f - cos_i, 5.f); }

inline float smith_ggx_correlated(float cos_o, float cos_i, float roughness)
{
    float a = roughness * roughness, a2 = a * a;
    float ggx_i = cos_o * std::sqrt((-cos_i * a2 + cos_i) * cos_i + a2);
    float ggx_o = cos_i * std::sqrt((-cos_o * a2 + cos_o) * cos_o + a2);
    return 2.f * cos_o * cos_i / (ggx_o + ggx_i);
}

inline float geometry_schlick_ggx(float cos_theta, float roughness)
{
    float a = roughness * roughness, a2 = a * a;
    float tan2 = (1.f - cos_theta * cos_theta) / (cos_theta * cos_theta);
    return 2.f / (1.f + std::sqrt(1.f + a2 * tan2));
}

// sampler::DistributionVn::Sample (core/math/Sampler.h:133-165): two draws, then the stretched-space VNDF construction
inline V3 sample_vndf(Rng &rng, V3 wo, float roughness)
{
    float a = roughness * roughness;
    float u1 = rng_unit(rng), u2 = rng_unit(rng);
    V3 v = normalized(mk(wo.x * a, wo.y * a, wo.z));
    V3 t1 = v.z < 0.999f ? normalized(cross(v, mk(0.f, 0.f, 1.f))) : mk(1.f, 0.f, 0.f);
    V3 t2 = cross(v, t1);
    float r = std::sqrt(u1);
    float phi = 2.f * kPi * u2;
    float p1 = r * m_cos(phi), p2 = r * m_sin(phi);
    float s = 0.5f * (1.f + v.z);
    p2 = (1.f - s) * std::sqrt(1.f - p1 * p1) + s * p2;
    float c = std::sqrt(std::max(0.f, 1.f - p1 * p1 - p2 * p2));
    V3 n = mk(p1 * t1.x + p2 * t2.x + c * v.x, p1 * t1.y + p2 * t2.y + c * v.y, p1 * t1.z + p2 * t2.z + c * v.z);
    return normalized(mk(a * n.x, a * n.y, std::max(0.f, n.z)));
}

// MaterialRenderProxy::SampleSurface: PbrMaterialRenderProxy.h:15-51 (SpecularBxDF then LambertianBxDF, BxDF.h:32-110)
// and DialetricMaterialRenderProxy.h:16-38 (DieletricBxDF, BxDF.h:112-141). Returns this bounce's throughput.
V3 sample_surface(const SceneView &sv, Rng &rng, const Surface &sf, V3 ray_dir, V3 &wi_world)
{
    const SpkSceneDesc *d = sv.desc;
    const SpkMaterialDesc &m = d->materials[sf.material];
    MaterialSample ms = material_sample(d, m, sf.u, sf.v);
    // utilities::GetLocalAxisFromNormal(normal, tangent) (Utilities.h:236-244)
    V3 au = normalized(cross(sf.normal, sf.tangent)), av = cross(sf.normal, au), aw = sf.normal;
    V3 w_o = -ray_dir;
    V3 lwo = mk(dot(w_o, au), dot(w_o, av), dot(w_o, aw)); // TransformBasisToLocal
    V3 lwi = mk(0.f, 0.f, 0.f);
    V3 throughput;
    if (m.type == SPK_MAT_DIELECTRIC)
    {
        float eta = m.eta;
        float cos_i = lwo.z, eta_i = 1.f, eta_t = eta; // FrDielectric(CosTheta(w_o), EtaI = 1, eta)
        if (!(cos_i > 0.f))
        {
            std::swap(eta_i, eta_t);
            cos_i = -cos_i;
        }
        float sin_i = std::sqrt(1 - cos_i * cos_i);
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
            fresnel = schlick(cos_i, r0);
        }
        if (rng_unit(rng) < fresnel)
        {
            lwi = mk(-lwo.x, -lwo.y, lwo.z); // utilities::Reflect(w) (Utilities.h:201-204)
        }
        else
        {
            bool entering = lwo.z > 0;
            float ratio = entering ? 1.0f / eta : eta / 1.0f;
            // utilities::Refract (Utilities.h:211-220)
            float ct = lwo.z;
            float sin_i2 = 1.f - ct * ct;
            float sin_t2 = ratio * ratio * sin_i2;
            float cos_t = std::sqrt(1 - sin_t2) * (ct > 0 ? -1.f : 1.f);
            lwi = normalized(mk(-ratio * lwo.x, -ratio * lwo.y, cos_t));
        }
        throughput = ms.base_color;
    }
    else
    {
        V3 wm = sample_vndf(rng, lwo, ms.roughness);
        float k = dot(lwo, wm) * 2.f;
        V3 swi = mk(wm.x * k - lwo.x, wm.y * k - lwo.y, wm.z * k - lwo.z); // utilities::Reflect(w_o, w_m)
        // SaturatedCosTheta = max(abs(z), 0)
        float cos_o = std::max(std::abs(lwo.z), 0.f), cos_in = std::max(std::abs(swi.z), 0.f);
        float cos_m = std::max(std::abs(wm.z), 0.f);
        float local_cos_i = dot(swi, wm);
        bool valid = false;
        V3 fresnel_color = mk(0.f, 0.f, 0.f);
        if (cos_o > kEps && cos_in > kEps && cos_m > kEps)
        {
            if (rng_unit(rng) < ms.metallic)
            {
                float f = m_pow(1.f - local_cos_i, 5.f); // SchlickApproximation(cos, Vector3 r0)
                fresnel_color = mk(ms.base_color.x + (1.f - ms.base_color.x) * f,
                                   ms.base_color.y + (1.f - ms.base_color.y) * f,
                                   ms.base_color.z + (1.f - ms.base_color.z) * f);
                valid = true;
            }
            else
            {
                float fresnel = schlick(local_cos_i, 0.04f); // F0
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
            // LambertianBxDF::Sample -> sampler::CosineWeightedHemiSphere::Sample (Sampler.h:55-66): unit disk lifted
            float dx, dy;
            unit_disk(rng, dx, dy);
            float z = std::sqrt(std::max(0.f, 1.0f - (dx * dx + dy * dy)));
            lwi = mk(dx, dy, z);
            throughput = ms.base_color;
        }
    }
    // utilities::TransformBasisToWorld(local_w_i, u, v, w).normalized()
    wi_world = normalized(mk(au.x * lwi.x + av.x * lwi.y + aw.x * lwi.z, au.y * lwi.x + av.y * lwi.y + aw.y * lwi.z,
                             au.z * lwi.x + av.z * lwi.y + aw.z * lwi.z));
    return throughput;
}

// ---------------------------------------------------------------------------------------------------------------------
// a14. SkyRenderProxy::Evaluate (renderer/proxy/SkyRenderProxy.cpp:50-64); Image2DCube::DirectionToTextureCoordinate and
// ::Sample (io/Image.cpp:298-357): major-axis face, uv = (+-a/b + 1) / 2, face-local bilinear
V3 sky_evaluate(const SpkSkyDesc &sky, V3 d)
{
    if (sky.has_map)
    {
        float ax = std::abs(d.x), ay = std::abs(d.y), az = std::abs(d.z), u, v;
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
        return sample_texture(sky.faces[face], u, v);
    }
    float t = 0.5f * d.z + 0.5f; // Lerp(Ones, color, t)
    return mk(1.f + (sky.color[0] - 1.f) * t, 1.f + (sky.color[1] - 1.f) * t, 1.f + (sky.color[2] - 1.f) * t);
}

// ---------------------------------------------------------------------------------------------------------------------
// a3-a5. RenderPixel / SetupViewRay / SamplePixel (renderer/renderer/CPURenderer.cpp:220-417), DebugMode::Color only
struct PixelSample
{
    V3 color = mk(0.f, 0.f, 0.f);
    V3 world_normal = mk(0.f, 0.f, 0.f);
    float valid = 1.f;
    uint32_t bounces = 0;
};

PixelSample render_pixel(const SceneView &sv, const SpkCamera &cam, uint32_t width, uint32_t height, uint32_t i,
                         uint32_t j, uint32_t frame_seed, uint32_t max_bounce, float output_limit, uint32_t mode,
                         Counters &cnt)
{
    const SpkSceneDesc *d = sv.desc;
    // BasePass (CPURenderer.cpp:419-450): pixel extents and the per-pixel reseed, unsigned 32-bit wrap
    const float pw = 1.f / static_cast<float>(width - 1), ph = 1.f / static_cast<float>(height - 1);
    Rng rng = rng_seed(j * width + i + frame_seed * width * height);
    // RenderPixel (:398-417): jitter u then v
    float u = (static_cast<float>(i) + rng_unit(rng)) * pw;
    float v = (static_cast<float>(j) + rng_unit(rng)) * ph;
    // SetupViewRay (:220-238): the lens disk is always drawn
    float dx, dy;
    unit_disk(rng, dx, dy);
    dx = dx * cam.aperture_radius;
    dy = dy * cam.aperture_radius;
    V3 origin = mk(cam.position) + (dx * mk(cam.right) + dy * mk(cam.up));
    V3 target = mk(cam.lower_left) + u * mk(cam.max_u) + v * mk(cam.max_v);
    V3 dir = normalized(target - origin);

    PixelSample r;
    V3 throughput = mk(1.f, 1.f, 1.f);
    uint32_t bounce = 0;
    for (; bounce < max_bounce; bounce++)
    {
        Candidate cand;
        trace_closest(sv, origin, dir, mode, cand, cnt);
        if (cand.prim == SPK_INVALID_ID) // terminal: hit nothing
        {
            if (d->sky.has_sky)
            {
                r.color = r.color + cmul(sky_evaluate(d->sky, dir), throughput);
            }
            break;
        }
        Surface sf;
        get_intersection(sv, origin, dir, cand, sf);
        if (bounce == 0)
        {
            r.world_normal = sf.normal;
        }
        V3 emissive = material_emissive(d, d->materials[sf.material], sf.u, sf.v);
        if (norm(emissive) > kEps) // terminal: emissive
        {
            r.color = r.color + cmul(emissive, throughput);
            break;
        }
        V3 next_dir;
        V3 this_throughput = sample_surface(sv, rng, sf, dir, next_dir);
        throughput = cmul(throughput, this_throughput);
        if (sqnorm(throughput) < kEps) // terminal: early out
        {
            r.valid = -1.f;
            break;
        }
        if (bounce >= 3) // Russian roulette
        {
            float p = std::max(throughput.x, std::max(throughput.y, throughput.z));
            p = std::clamp(p, 0.05f, 1.0f);
            if (rng_unit(rng) > p)
            {
                break;
            }
            throughput = throughput / p;
        }
        origin = sf.location + next_dir * kTolerance;
        dir = next_dir;
    }
    r.bounces = bounce;
    r.color = mk(std::min(r.color.x, output_limit), std::min(r.color.y, output_limit), std::min(r.color.z, output_limit));
    cnt.rng_draws += rng.draws;
    return r;
}

// row-parallel worker pool: one task per image row like BasePass (CPURenderer.cpp:435-447); rows are independent
template <typename RowFn> void parallel_rows(uint32_t height, uint32_t threads, RowFn &&fn)
{
    uint32_t nt = std::max(1u, threads);
    if (nt == 1)
    {
        for (uint32_t j = 0; j < height; j++)
        {
            fn(j, 0u);
        }
        return;
    }
    std::atomic<uint32_t> next{0};
    std::vector<std::thread> pool;
    for (uint32_t t = 0; t < nt; t++)
    {
        pool.emplace_back([&, t]() {
            for (;;)
            {
                uint32_t j = next.fetch_add(1);
                if (j >= height)
                {
                    break;
                }
                fn(j, t);
            }
        });
    }
    for (auto &th : pool)
    {
        th.join();
    }
}

// ACESFilm (CPURenderer.cpp:554-563)
inline float aces(float c)
{
    float r = (c * (c * 2.51f + 0.03f)) / (c * (c * 2.43f + 0.59f) + 0.14f);
    return std::min(std::max(r, 0.f), 1.f);
}

// scene view cache: rebuilt when the descriptor's identity changes (the tests reuse one descriptor for many calls)
struct CachedView
{
    const SpkSceneDesc *key = nullptr;
    uint64_t stamp = 0;
    SceneView view;
};
CachedView g_cache;

uint64_t scene_stamp(const SpkSceneDesc *d)
{
    uint64_t h = 1469598103934665603ull;
    auto mix = [&h](uint64_t v) { h = (h ^ v) * 1099511628211ull; };
    mix(d->num_meshes);
    mix(d->num_primitives);
    mix(d->num_materials);
    mix(reinterpret_cast<uint64_t>(d->primitives));
    mix(reinterpret_cast<uint64_t>(d->meshes));
    mix(reinterpret_cast<uint64_t>(d->tlas.nodes));
    // addresses are not an identity (freed tables are handed out again): the heads of the cached tables by content too
    auto bytes = [&](const void *p, size_t n) {
        const auto *b = static_cast<const uint8_t *>(p);
        for (size_t k = 0; p && k < n; k++)
        {
            mix(b[k]);
        }
    };
    auto tree = [&](const SpkBvhDesc &t) {
        mix(t.num_nodes);
        mix(t.num_prims);
        bytes(t.nodes, size_t(std::min<uint32_t>(t.num_nodes, 16u)) * t.node_stride);
        bytes(t.prim_ids, size_t(std::min<uint32_t>(t.num_prims, 64u)) * t.prim_id_bytes);
    };
    tree(d->tlas);
    for (uint32_t m = 0; m < d->num_meshes; m++)
    {
        mix(reinterpret_cast<uint64_t>(d->meshes[m].vertices));
        mix(d->meshes[m].num_faces);
        mix(d->meshes[m].num_vertices);
        bytes(d->meshes[m].vertices, size_t(std::min<uint32_t>(d->meshes[m].num_vertices, 32u)) * 12);
        bytes(d->meshes[m].indices, size_t(std::min<uint32_t>(d->meshes[m].num_faces, 32u)) * 12);
        tree(d->meshes[m].blas);
    }
    return h;
}

const SceneView &view_of(const SpkSceneDesc *d)
{
    uint64_t stamp = scene_stamp(d);
    if (g_cache.key != d || g_cache.stamp != stamp)
    {
        g_cache.view = SceneView{};
        build_view(d, g_cache.view);
        g_cache.key = d;
        g_cache.stamp = stamp;
    }
    return g_cache.view;
}

inline SpkHit to_hit(const Candidate &c)
{
    SpkHit h;
    h.t = c.t;
    h.u = c.u;
    h.v = c.v;
    h.prim = c.prim;
    h.face = c.face;
    if (c.prim == SPK_INVALID_ID)
    {
        h.u = 0.f;
        h.v = 0.f;
        h.face = SPK_INVALID_ID;
    }
    return h;
}
} // namespace

extern "C"
{
const char *orc_kind(void) { return "port"; }

int orc_set_libm(uint32_t mode)
{
    g_libm_mode = mode ? 1 : 0;
    return 0;
}

void orc_rng_state(uint32_t seed, uint32_t state_out[4])
{
    Rng r = rng_seed(seed);
    memcpy(state_out, r.s, 16);
}

void orc_rng_units(uint32_t seed, uint32_t n, float *out)
{
    Rng r = rng_seed(seed);
    for (uint32_t k = 0; k < n; k++)
    {
        out[k] = rng_unit(r);
    }
}

void orc_kat_vec(uint32_t op, const float *a, const float *b, const float *m, float *out)
{
    V3 va = mk(a), vb = b ? mk(b) : mk(0.f, 0.f, 0.f), r = mk(0.f, 0.f, 0.f);
    const float identity[12] = {1, 0, 0, 0, 0, 1, 0, 0, 0, 0, 1, 0};
    const float *mm = m ? m : identity;
    switch (op)
    {
    case 0:
        out[0] = dot(va, vb);
        return;
    case 1:
        out[0] = sqnorm(va);
        return;
    case 2:
        r = normalized(va);
        break;
    case 3:
        r = cross(va, vb);
        break;
    case 4:
        r = xform_dir(mm, va);
        break;
    case 5:
        r = xform_point(mm, va);
        break;
    default:
        break;
    }
    out[0] = r.x;
    out[1] = r.y;
    out[2] = r.z;
}

float orc_kat_libm(uint32_t op, float x, float y)
{
    switch (op)
    {
    case 0:
        return std::sin(x);
    case 1:
        return std::cos(x);
    case 2:
        return std::pow(x, y);
    case 3:
        return std::sqrt(x);
    case 4:
        return std::acos(x);
    case 5:
        return std::tan(x);
    case 6:
        return std::atan2(x, y);
    default:
        return 0.f;
    }
}

void orc_kat_unit_disk(uint32_t seed, float *out)
{
    Rng r = rng_seed(seed);
    unit_disk(r, out[0], out[1]);
}

void orc_kat_texture(const SpkTextureDesc *tex, float u, float v, float *out)
{
    V3 c = sample_texture(*tex, u, v);
    out[0] = c.x;
    out[1] = c.y;
    out[2] = c.z;
}

void orc_kat_sky(const SpkSkyDesc *sky, const float *direction, float *out)
{
    V3 c = sky->has_sky ? sky_evaluate(*sky, mk(direction)) : mk(0.f, 0.f, 0.f);
    out[0] = c.x;
    out[1] = c.y;
    out[2] = c.z;
}

uint32_t orc_kat_sample_surface(const SpkSceneDesc *scene, uint32_t material, uint32_t seed, const float *dir,
                                const float *normal, const float *tangent, const float *uv, float *out_wi,
                                float *out_throughput)
{
    SceneView sv; // materials and textures are read straight from the descriptor: no BVH view needed
    sv.desc = scene;
    Surface sf;
    sf.location = mk(0.f, 0.f, 0.f);
    sf.normal = mk(normal);
    sf.tangent = mk(tangent);
    sf.u = uv[0];
    sf.v = uv[1];
    sf.material = material;
    Rng rng = rng_seed(seed);
    V3 wi = mk(0.f, 0.f, 0.f);
    V3 thr = sample_surface(sv, rng, sf, mk(dir), wi);
    out_wi[0] = wi.x;
    out_wi[1] = wi.y;
    out_wi[2] = wi.z;
    out_throughput[0] = thr.x;
    out_throughput[1] = thr.y;
    out_throughput[2] = thr.z;
    return static_cast<uint32_t>(rng.draws);
}

int orc_trace_closest(const SpkSceneDesc *scene, const SpkRay *rays, uint64_t num_rays, SpkHit *hits, uint32_t mode,
                      uint32_t threads, OrcCounters *counters)
{
    if (!scene || (!rays && num_rays) || (!hits && num_rays) || mode > ORC_TRACE_BVH_CULLED)
    {
        return SPK_ERR_INVALID_ARGUMENT;
    }
    const SceneView &sv = view_of(scene);
    uint32_t nt = std::max(1u, threads);
    if (num_rays < 1024)
    {
        nt = 1;
    }
    std::vector<Counters> per_thread(nt);
    const uint64_t chunk = 4096;
    uint32_t chunks = static_cast<uint32_t>((num_rays + chunk - 1) / chunk);
    parallel_rows(chunks, nt, [&](uint32_t c, uint32_t t) {
        uint64_t b = uint64_t(c) * chunk, e = std::min(num_rays, b + chunk);
        for (uint64_t k = b; k < e; k++)
        {
            Candidate cand;
            trace_closest(sv, mk(rays[k].origin), mk(rays[k].direction), mode, cand, per_thread[t]);
            hits[k] = to_hit(cand);
        }
    });
    if (counters)
    {
        *counters = OrcCounters{};
        for (const auto &c : per_thread)
        {
            c.AddTo(*counters);
        }
    }
    return SPK_OK;
}

uint32_t orc_intersect_full(const SpkSceneDesc *scene, const SpkRay *ray, float *out12)
{
    const SceneView &sv = view_of(scene);
    Counters cnt;
    Candidate cand;
    V3 o = mk(ray->origin), dvec = mk(ray->direction);
    trace_closest(sv, o, dvec, ORC_TRACE_BVH_REFERENCE, cand, cnt);
    if (cand.prim == SPK_INVALID_ID)
    {
        return SPK_INVALID_ID;
    }
    Surface sf;
    get_intersection(sv, o, dvec, cand, sf);
    out12[0] = cand.t;
    out12[1] = sf.location.x;
    out12[2] = sf.location.y;
    out12[3] = sf.location.z;
    out12[4] = sf.normal.x;
    out12[5] = sf.normal.y;
    out12[6] = sf.normal.z;
    out12[7] = sf.tangent.x;
    out12[8] = sf.tangent.y;
    out12[9] = sf.tangent.z;
    out12[10] = sf.u;
    out12[11] = sf.v;
    return cand.prim;
}

int orc_render(const SpkSceneDesc *scene, const SpkCamera *camera, uint32_t width, uint32_t height,
               const SpkRenderParams *params, float *accum, uint32_t *cumulated, uint32_t trace_mode, uint32_t threads,
               OrcCounters *counters)
{
    if (!scene || !camera || !params || !accum || !cumulated || trace_mode > ORC_TRACE_BVH_CULLED || width < 2 ||
        height < 2)
    {
        return SPK_ERR_INVALID_ARGUMENT;
    }
    const SceneView &sv = view_of(scene);
    uint32_t nt = std::max(1u, threads);
    std::vector<Counters> per_thread(nt);
    std::vector<float> frame(size_t(width) * height * 4); // gbuffer_.color of one sample
    const bool spatial = params->flags & SPK_FLAG_SPATIAL_DENOISE;
    if (spatial && params->accum_mode != SPK_ACCUM_RUNNING_AVERAGE)
    {
        return SPK_ERR_UNSUPPORTED; // the reference's DenoisePass is the running average
    }
    std::vector<float> normals(spatial ? size_t(width) * height * 3 : 0); // gbuffer_.world_normal
    std::vector<float> filled;                                             // ping_pong_buffer_
    uint32_t stride = std::max(1u, params->sample_stride);
    for (uint32_t s = 0; s < params->sample_count; s++)
    {
        uint32_t frame_seed = params->first_sample + s * stride; // dispatched_sample_count_ (CPURenderer.cpp:432)
        parallel_rows(height, nt, [&](uint32_t j, uint32_t t) {
            for (uint32_t i = 0; i < width; i++)
            {
                PixelSample ps = render_pixel(sv, *camera, width, height, i, j, frame_seed, params->max_bounce,
                                              params->output_limit, trace_mode, per_thread[t]);
                float *px = frame.data() + (size_t(j) * width + i) * 4;
                px[0] = ps.color.x;
                px[1] = ps.color.y;
                px[2] = ps.color.z;
                px[3] = ps.valid;
                if (spatial)
                {
                    float *pn = normals.data() + (size_t(j) * width + i) * 3;
                    pn[0] = ps.world_normal.x;
                    pn[1] = ps.world_normal.y;
                    pn[2] = ps.world_normal.z;
                }
            }
        });
        if (spatial)
        {
            // DenoisePass, spatial part (CPURenderer.cpp:511-526) + SpatialDenoisePixel (:452-504), rows in order (one
            // worker): each valid non-sky pixel gives its colour to the FIRST of its 8 neighbours, in the reference's
            // direction order, that is invalid and has normal similarity >= 0.99; later writers overwrite earlier ones
            static const int dirs[8][2] = {{1, 0}, {0, 1}, {-1, 0}, {0, -1}, {-1, -1}, {1, -1}, {-1, 1}, {1, 1}};
            filled = frame;
            for (uint32_t j = 0; j < height; j++)
            {
                for (uint32_t i = 0; i < width; i++)
                {
                    const float *c = frame.data() + (size_t(j) * width + i) * 4;
                    const float *n = normals.data() + (size_t(j) * width + i) * 3;
                    if (!(c[3] > 0.f) || (n[0] == 0.f && n[1] == 0.f && n[2] == 0.f))
                    {
                        continue;
                    }
                    for (int k = 0; k < 8; k++)
                    {
                        int si = int(i) + dirs[k][0], sj = int(j) + dirs[k][1];
                        if (si < 0 || si >= int(width) || sj < 0 || sj >= int(height))
                        {
                            continue;
                        }
                        const float *nc = frame.data() + (size_t(sj) * width + si) * 4;
                        const float *nn = normals.data() + (size_t(sj) * width + si) * 3;
                        if (nc[3] > 0.f)
                        {
                            continue;
                        }
                        float similarity = (nn[0] * n[0] + nn[1] * n[1]) + nn[2] * n[2];
                        if (similarity < 0.99f)
                        {
                            continue;
                        }
                        float *o = filled.data() + (size_t(sj) * width + si) * 4;
                        o[0] = c[0];
                        o[1] = c[1];
                        o[2] = c[2];
                        o[3] = 1.f;
                        break;
                    }
                }
            }
            frame.swap(filled);
        }
        if (params->accum_mode == SPK_ACCUM_RUNNING_AVERAGE)
        {
            // DenoisePass, temporal part (CPURenderer.cpp:526-544): acc = Lerp(new, acc, N / (N + spp)), spp = 1
            // CameraRenderProxy::Update caps the cumulated count at max_spp (CameraRenderProxy.cpp:70-73)
            const uint32_t capped = params->cumulated_cap ? std::min(*cumulated, params->cumulated_cap) : *cumulated;
            float ma = static_cast<float>(capped) / static_cast<float>(capped + 1u);
            for (size_t k = 0; k < frame.size(); k++)
            {
                accum[k] = lerp1(frame[k], accum[k], ma);
            }
        }
        else
        {
            for (size_t k = 0; k < frame.size(); k++)
            {
                accum[k] += frame[k];
            }
        }
        *cumulated += 1;
    }
    if (counters)
    {
        *counters = OrcCounters{};
        counters->paths = uint64_t(width) * height * params->sample_count;
        for (const auto &c : per_thread)
        {
            c.AddTo(*counters);
        }
    }
    return SPK_OK;
}

int orc_sample_pixel(const SpkSceneDesc *scene, const SpkCamera *camera, uint32_t width, uint32_t height, uint32_t i,
                     uint32_t j, uint32_t frame_seed, uint32_t max_bounce, float output_limit, uint32_t trace_mode,
                     float *out8, OrcCounters *counters)
{
    if (!scene || !camera || !out8 || trace_mode > ORC_TRACE_BVH_CULLED)
    {
        return SPK_ERR_INVALID_ARGUMENT;
    }
    const SceneView &sv = view_of(scene);
    Counters cnt;
    PixelSample ps = render_pixel(sv, *camera, width, height, i, j, frame_seed, max_bounce, output_limit, trace_mode, cnt);
    out8[0] = ps.color.x;
    out8[1] = ps.color.y;
    out8[2] = ps.color.z;
    out8[3] = ps.valid;
    out8[4] = ps.world_normal.x;
    out8[5] = ps.world_normal.y;
    out8[6] = ps.world_normal.z;
    out8[7] = static_cast<float>(ps.bounces);
    if (counters)
    {
        *counters = OrcCounters{};
        counters->paths = 1;
        cnt.AddTo(*counters);
    }
    return SPK_OK;
}

// a16. ToneMappingPass (CPURenderer.cpp:565-576) + Image2D::SetPixel (io/Image.h:154-184): ACES per channel, row H-1-j;
// RGBAFloat16 stores Eigen::half(rgb), alpha 1; an sRGB8 image stores LinearToSrgb = pow(0.45) then (v * 255) truncated
int orc_resolve(const float *accum, uint32_t width, uint32_t height, uint32_t accum_mode, uint32_t format,
                uint32_t total_samples, float exposure, void *out)
{
    if (!accum || !out || format > SPK_RESOLVE_TONEMAP_SRGB8)
    {
        return SPK_ERR_INVALID_ARGUMENT;
    }
    const float div = accum_mode == SPK_ACCUM_SUM ? static_cast<float>(std::max(1u, total_samples)) : 1.f;
    for (uint32_t j = 0; j < height; j++)
    {
        for (uint32_t i = 0; i < width; i++)
        {
            const float *p = accum + (size_t(j) * width + i) * 4;
            float c[4] = {p[0], p[1], p[2], p[3]};
            if (accum_mode == SPK_ACCUM_SUM)
            {
                for (float &x : c)
                {
                    x = x / div;
                }
            }
            if (format == SPK_RESOLVE_LINEAR_F32)
            {
                memcpy(static_cast<float *>(out) + (size_t(j) * width + i) * 4, c, 16);
                continue;
            }
            float r = aces(c[0] * exposure), g = aces(c[1] * exposure), b = aces(c[2] * exposure);
            size_t o = size_t(height - 1 - j) * width + i;
            if (format == SPK_RESOLVE_TONEMAP_F16)
            {
                uint16_t *h = static_cast<uint16_t *>(out) + o * 4;
                h[0] = float_to_half(r);
                h[1] = float_to_half(g);
                h[2] = float_to_half(b);
                h[3] = float_to_half(1.f);
            }
            else
            {
                uint8_t *q = static_cast<uint8_t *>(out) + o * 4;
                q[0] = static_cast<uint8_t>(m_pow(r, 0.45f) * 255.f);
                q[1] = static_cast<uint8_t>(m_pow(g, 0.45f) * 255.f);
                q[2] = static_cast<uint8_t>(m_pow(b, 0.45f) * 255.f);
                q[3] = 255;
            }
        }
    }
    return SPK_OK;
}
} // extern "C"
