/*
 * spk_device.cuh — device-side data layout and the bit-exact arithmetic vocabulary of the `cuda` pipeline.
 *
 * Layout in HBM (all arrays 16-byte aligned, built once by spk_cuda_upload_scene, read-only afterwards):
 *   node pairs   64 B per inner-node visit: the two adjacent children nodes[first_id], nodes[first_id+1] of the
 *                bvh::v2 layout (core/math/BVH.h:18-19) as 2 x {bounds[6], index, pad}. Because the library places
 *                children at odd/even ids (1,2),(3,4),.. the device array is nodes[1..] so every pair is 64-B aligned.
 *                TLAS and all BLAS pair arrays live in ONE buffer; an inner index word is rewritten at upload to the
 *                absolute pair slot (slot<<4, count 0) so traversal needs no per-mesh base.
 *   triangles    48 B per triangle, stored in LEAF ORDER (bvh_.prim_ids order) so a leaf is one contiguous run:
 *                {p0.xyz, face_idx bits}, {e1.xyz, 0}, {e2.xyz, 0} — e1/e2 are the host's floats
 *                (MeshRenderProxy.cpp:78-92), never recomputed on the device. A leaf index word is rewritten to the
 *                absolute triangle slot.
 *   instances    128 B per primitive in TLAS leaf order (SceneRenderProxy.cpp:59-68 permutes primitives_ the same
 *                way): inv 3x4, fwd 3x4 (bit copies of Transform::GetMatrix of the inverse / forward), root index
 *                word of the BLAS, kind, radius2, primitive index (SceneRenderProxy order), material, mesh.
 *
 * Arithmetic: this header is compiled with -fmad=false. Every reduction mirrors the association the reference's
 * Eigen expressions compile to with g++ -O2 -ffp-contract=off (re-probed against oracle/_ref):
 *   a.dot(b)        = (a0*b0 + a1*b1) + a2*b2
 *   a.squaredNorm() = a0*a0 + (a1*a1 + a2*a2)
 *   a.normalized()  = a / sqrt(squaredNorm)      (true division)
 *   M3x3 * v  (row) = m0*v0 + (m1*v1 + m2*v2)    (Transform::TransformDirection)
 *   Affine * p(row) = ((m0*p0 + m1*p1) + m2*p2) + t   (Transform::TransformPoint)
 *   cross           = textbook (a1*b2 - a2*b1, ...)
 */
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

#define SPK_INVALID_ID 0xFFFFFFFFu
#define SPK_FLT_MAX 3.402823466e+38f

namespace spk
{
struct __align__(16) DInstance
{
    float inv[12]; // rows 0..2 of the inverse transform
    float fwd[12]; // rows 0..2 of the forward transform
    uint32_t root; // BLAS root index word (absolute: slot<<4 | count), unused for spheres
    uint32_t kind; // 0 mesh, 1 sphere
    float radius2;
    uint32_t prim; // primitive index in SceneRenderProxy::primitives_ registration order
    uint32_t material;
    uint32_t mesh;
    uint32_t shade_class; // 1 + mesh + 2 dielectric + 4 textured + 8 emissive: the key segments are sorted by before shading
    uint32_t pad;
};
static_assert(sizeof(DInstance) == 128, "instance record");

struct __align__(16) DMesh
{
    const float *normals;    // 3 per vertex
    const float *tangents;   // 4 per vertex
    const float *uvs;        // 2 per vertex or nullptr
    const uint32_t *indices; // 3 per face
};

struct __align__(16) DMaterial
{
    uint32_t base_color_tex, emissive_tex, metallic_roughness_tex, normal_tex;
    float metallic, roughness, eta;
    uint32_t type;
    float base_color[3];
    float pad0;
    float emissive[3];
    float pad1;
};
static_assert(sizeof(DMaterial) == 64, "material record");

struct __align__(16) DTexture
{
    const void *pixels;
    uint32_t width, height;
    uint32_t format; // SpkPixelFormat
    uint32_t pad;
};

struct DSky
{
    uint32_t has_sky, has_map;
    float color[3];
    // the six faces of a sky map are entries face_base .. face_base + 5 of the texture table: a face picked by a run-time
    // index out of the kernel PARAMETERS made every thread of every kernel that takes a DScene copy the whole struct to
    // local memory first (wf_shade: 370 B of STL per thread, 4 GB of DRAM write-backs per launch of 16 Mi slots)
    uint32_t face_base;
};

struct DCamera
{
    float position[3], up[3], front[3], right[3], max_u[3], max_v[3], lower_left[3];
    float aperture_radius, exposure;
};

struct DScene
{
    const float4 *pairs;        // 4 float4 per pair
    const float4 *tris;         // 3 float4 per triangle
    const DInstance *instances; // TLAS leaf order
    const DMesh *meshes;
    const DMaterial *materials;
    const DTexture *textures;
    uint32_t tlas_root;         // index word of the TLAS root (absolute)
    uint32_t num_instances;     // 0 => empty scene
    float cull_extent;          // world extent used for the t-culling slack
    uint32_t top_base;          // first pair slot of the run a kernel may stage in shared memory (the hottest tree's top)
    uint32_t top_count;         // pairs in that run (level-ordered: any prefix is the top of the tree)
    uint32_t pad;
    DSky sky;
};

struct V3
{
    float x, y, z;
};

__device__ __forceinline__ V3 mk(float x, float y, float z)
{
    V3 r;
    r.x = x;
    r.y = y;
    r.z = z;
    return r;
}
__device__ __forceinline__ V3 operator+(V3 a, V3 b) { return mk(a.x + b.x, a.y + b.y, a.z + b.z); }
__device__ __forceinline__ V3 operator-(V3 a, V3 b) { return mk(a.x - b.x, a.y - b.y, a.z - b.z); }
__device__ __forceinline__ V3 operator-(V3 a) { return mk(-a.x, -a.y, -a.z); }
__device__ __forceinline__ V3 operator*(V3 a, float s) { return mk(a.x * s, a.y * s, a.z * s); }
__device__ __forceinline__ V3 operator*(float s, V3 a) { return mk(s * a.x, s * a.y, s * a.z); }
#ifdef SPK_NOINLINE_DIV3
static __device__ __noinline__ V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
#else
__device__ __forceinline__ V3 operator/(V3 a, float s) { return mk(a.x / s, a.y / s, a.z / s); }
#endif
__device__ __forceinline__ V3 cmul(V3 a, V3 b) { return mk(a.x * b.x, a.y * b.y, a.z * b.z); }
__device__ __forceinline__ float dot(V3 a, V3 b) { return (a.x * b.x + a.y * b.y) + a.z * b.z; }
__device__ __forceinline__ float sqnorm(V3 a) { return a.x * a.x + (a.y * a.y + a.z * a.z); }
// sqrtf(0) and 0/x leave the inlined fast paths of CUDA's IEEE sqrt and division (FCHK flags a zero operand) and call
// a ~50-instruction subroutine at the far end of the kernel: ncu showed 1.9 M such calls per 16-spp frame from
// norm(emissive = 0) and 1.3-1.9 M per component from normalizing axis-aligned vectors (floor normal (0,0,1)), each an
// instruction-cache miss. The guards below return the same bits (sqrt(+0) = +0; (+-0)/n = +-0 for n > 0) without it.
__device__ __forceinline__ float norm(V3 a)
{
    float sq = sqnorm(a);
    return sq > 0.f ? sqrtf(sq) : sq; // sq is +0 or NaN otherwise: sqrt(+0) = +0, sqrt(NaN) = NaN
}
__device__ __forceinline__ float div_by_norm(float a, float n, bool n_pos)
{
    const bool zero = a == 0.f && n_pos;
    float q = (zero ? n : a) / n;
    return zero ? a : q;
}
// ONE out-of-line copy instead of ~15 inlined ones (each a square root and three IEEE divisions with their slow-path
// stubs, ~60 instructions): the sorted megakernel shrank from 4880 to 4144 SASS instructions (76 -> 65 KB), and with it the
// instruction-cache misses that made small CTAs lose: TestScene 1080p x 32 spp 19.04 -> 18.04 ms with 16 warps x 2 CTAs
// and 21.69 -> 17.91 ms with 8 warps x 4 (B200; room 122.0 -> 120.8). Going further (rng_seed, make_slab_ray, V3 / float
// out of line as well) measured 0.5-4 % SLOWER each: those have one or two call sites and the call costs more than the
// bytes. -DSPK_INLINE_NORM restores the inlined form for A/B builds.
#ifndef SPK_INLINE_NORM
static __device__ __noinline__ V3 normalized(V3 a)
#else
__device__ __forceinline__ V3 normalized(V3 a)
#endif
{
    float sq = sqnorm(a);
    float n = sq > 0.f ? sqrtf(sq) : sq;
    const bool n_pos = n > 0.f; // false for the zero vector (0/0 = NaN must stay NaN) and for NaN
    return mk(div_by_norm(a.x, n, n_pos), div_by_norm(a.y, n, n_pos), div_by_norm(a.z, n, n_pos));
}
__device__ __forceinline__ V3 cross(V3 a, V3 b)
{
    return mk(a.y * b.z - a.z * b.y, a.z * b.x - a.x * b.z, a.x * b.y - a.y * b.x);
}
// Transform::TransformDirection: rows of the 3x3 block, m0*v0 + (m1*v1 + m2*v2)
__device__ __forceinline__ V3 xform_dir(const float *m, V3 v)
{
    return mk(m[0] * v.x + (m[1] * v.y + m[2] * v.z), m[4] * v.x + (m[5] * v.y + m[6] * v.z),
              m[8] * v.x + (m[9] * v.y + m[10] * v.z));
}
// Transform::TransformPoint: ((m0*p0 + m1*p1) + m2*p2) + t
__device__ __forceinline__ V3 xform_point(const float *m, V3 p)
{
    return mk(((m[0] * p.x + m[1] * p.y) + m[2] * p.z) + m[3], ((m[4] * p.x + m[5] * p.y) + m[6] * p.z) + m[7],
              ((m[8] * p.x + m[9] * p.y) + m[10] * p.z) + m[11]);
}

// bvh::v2 safe_inverse: 1/x, or copysign(FLT_MAX, x) when |x| <= FLT_EPSILON
__device__ __forceinline__ float safe_inverse(float x)
{
    return fabsf(x) <= 1.1920929e-07f ? copysignf(SPK_FLT_MAX, x) : 1.f / x;
}

// One 256-bit read-only load (LDG.E.256 on sm_100a; PTX ld.global.nc.v8.f32, 32-byte aligned address). A 64-byte node
// pair is two of these instead of four LDG.128: incoherent lanes cost one L1 wavefront per lane per load instruction,
// and the ray-batch kernel sat at 0.88 wavefronts per clock per SM (L1/TEX 79-86 % busy) whatever the scheduling.
__device__ __forceinline__ void ldg256(const float4 *p, float4 &a, float4 &b)
{
    asm("ld.global.nc.v8.f32 {%0,%1,%2,%3,%4,%5,%6,%7}, [%8];"
                 : "=f"(a.x), "=f"(a.y), "=f"(a.z), "=f"(a.w), "=f"(b.x), "=f"(b.y), "=f"(b.z), "=f"(b.w)
                 : "l"(p));
}

struct Hit
{
    float t, u, v;
    uint32_t inst; // instance slot (TLAS leaf order) or SPK_INVALID_ID
    uint32_t face;
};
} // namespace spk
