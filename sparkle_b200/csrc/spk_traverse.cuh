/*
 * spk_traverse.cuh — software BVH traversal (B200 has no RT cores), shared by the ray-batch kernel and the path tracer.
 *
 * Follows, operation for operation:
 *   TLAS::IntersectInternal<false>   libraries/source/renderer/proxy/SceneRenderProxy.cpp:82-126
 *   BLAS::IntersectInternal<false>   libraries/source/renderer/proxy/MeshRenderProxy.cpp:94-126
 *   BLAS::IntersectTriangle<false>   libraries/source/renderer/proxy/MeshRenderProxy.cpp:128-194 (Moeller-Trumbore, not
 *                                    watertight: reproducing it is what makes closest-hit ids and t bit-exact)
 *   SphereRenderProxy::IntersectInternal<false>  libraries/source/renderer/proxy/SphereRenderProxy.cpp:5-38
 *   bvh::v2::Bvh::intersect<false,false> (madmann91/bvh @5a9d759c, absent from the mount; published algorithm): safe
 *                                    inverse direction, slab test as separate multiply and add, near child first, far
 *                                    child pushed, no swap on equal entry distance.
 *
 * CULL=false reproduces the reference exactly (its bvh_ray.tmax is never shortened). CULL=true skips a child
 * whose entry distance exceeds best_t + slack; the children that remain are visited in the reference order, so the
 * first-found winner among equal t is unchanged. slack covers the difference between the object-space parameter the
 * boxes see and the world_t = (p.x - o.x)/d.x the reference compares (Ray.h:43-47): slack = best_t*2^-12 +
 * extent*2^-16/|d.x| (grows without bound as d.x -> 0, where world_t itself is noise: then nothing is culled).
 */
#pragma once
#include "spk_device.cuh"

namespace spk
{
struct TraceCounters
{
    uint32_t node_visits, tri_tests, inst_visits, sphere_tests;
};

struct SlabRay
{
    float ix, iy, iz; // safe inverse direction
    float ox, oy, oz; // -origin * inverse direction
    bool nx, ny, nz;  // octant: direction sign bits
};

#ifdef SPK_NOINLINE_SLAB
static __device__ __noinline__ SlabRay make_slab_ray(V3 o, V3 d)
#else
__device__ __forceinline__ SlabRay make_slab_ray(V3 o, V3 d)
#endif
{
    SlabRay r;
    r.ix = safe_inverse(d.x);
    r.iy = safe_inverse(d.y);
    r.iz = safe_inverse(d.z);
    r.ox = -o.x * r.ix;
    r.oy = -o.y * r.iy;
    r.oz = -o.z * r.iz;
    r.nx = signbit(d.x);
    r.ny = signbit(d.y);
    r.nz = signbit(d.z);
    return r;
}

// Node::intersect_fast: returns entry distance t0 and sets hit = (t0 <= t1); ray.tmin = 0, ray.tmax = FLT_MAX
__device__ __forceinline__ float slab(const SlabRay &r, float minx, float maxx, float miny, float maxy, float minz,
                                      float maxz, bool &hit)
{
    float lox = (r.nx ? maxx : minx) * r.ix + r.ox;
    float hix = (r.nx ? minx : maxx) * r.ix + r.ox;
    float loy = (r.ny ? maxy : miny) * r.iy + r.oy;
    float hiy = (r.ny ? miny : maxy) * r.iy + r.oy;
    float loz = (r.nz ? maxz : minz) * r.iz + r.oz;
    float hiz = (r.nz ? minz : maxz) * r.iz + r.oz;
    // robust_max(a,b) = a > b ? a : b drops a NaN first operand, as fmaxf does
    float t0 = fmaxf(loz, fmaxf(loy, fmaxf(lox, 0.f)));
    float t1 = fminf(hiz, fminf(hiy, fminf(hix, SPK_FLT_MAX)));
    hit = t0 <= t1;
    return t0;
}

// BLAS::IntersectTriangle<false>. o/d: local ray; wo/wd: world ray; fwd: instance forward matrix rows.
__device__ __forceinline__ void intersect_triangle(const float4 *__restrict__ tri, V3 o, V3 d, V3 wo, V3 wd,
                                                   const float *fwd, uint32_t inst, Hit &best)
{
    float4 a0 = __ldg(tri), a1 = __ldg(tri + 1), a2 = __ldg(tri + 2);
    V3 p0 = mk(a0.x, a0.y, a0.z), e1 = mk(a1.x, a1.y, a1.z), e2 = mk(a2.x, a2.y, a2.z);
    V3 h = cross(d, e2);
    float a = dot(e1, h);
    if (fabsf(a) < 1e-6f)
    {
        return;
    }
    V3 s = o - p0;
    float f = 1.f / a;
    float u = f * dot(s, h);
    if (u < 0.f || u > 1.f)
    {
        return;
    }
    V3 q = cross(s, e1);
    float v = f * dot(d, q);
    if (v < 0.f || u + v > 1.f)
    {
        return;
    }
    float t = f * dot(e2, q);
    if (t > 1e-6f)
    {
        V3 lp = o + d * t;                    // Ray::At
        // Transform::TransformPoint, then Ray::InverseAt which reads the x component only: row 0 suffices
        float wpx = ((fwd[0] * lp.x + fwd[1] * lp.y) + fwd[2] * lp.z) + fwd[3];
        float world_t = (wpx - wo.x) / wd.x;
        if (world_t > 0.f && world_t < best.t)
        {
            best.t = world_t;
            best.u = u;
            best.v = v;
            best.face = __float_as_uint(a0.w);
            best.inst = inst;
        }
    }
}

// SphereRenderProxy::IntersectInternal<false>
__device__ __forceinline__ void intersect_sphere(const DInstance *__restrict__ in, V3 wo, V3 wd, uint32_t inst,
                                                 Hit &best)
{
    V3 center = mk(__ldg(&in->fwd[3]), __ldg(&in->fwd[7]), __ldg(&in->fwd[11]));
    V3 oc = wo - center;
    float half_b = dot(wd, oc);
    float c = sqnorm(oc) - __ldg(&in->radius2);
    float disc = half_b * half_b - 1.f * c;
    if (disc < 0.f)
    {
        return;
    }
    float sq = sqrtf(disc);
    float t = (-half_b + (c >= 0.f ? -sq : sq)) / 1.f;
    if (t > 0.f && t < best.t)
    {
        best.t = t;
        best.u = 0.f; // SphereRenderProxy::GetIntersection ignores the candidate's u, v (SphereRenderProxy.h:22-31)
        best.v = 0.f;
        best.inst = inst;
        best.face = SPK_INVALID_ID;
    }
}

constexpr uint32_t kLeaveInstance = 0xFFFFFFFFu; // stack marker: the BLAS of the current instance is exhausted
constexpr uint32_t kTraversalDone = 0xFFFFFFFEu; // stack marker under everything: nothing left
constexpr int kThreadStack = 76;                  // 32 TLAS + 1 leaf remainder + 6 world slab words + 1 marker + 32 BLAS (+2)

// SceneRenderProxy::Intersect<false> up to the winning candidate (GetIntersection is done by the shading stage).
// ONE loop for both levels (the reference nests BLAS::Intersect inside the TLAS leaf callback; written that way the
// lanes of a warp wait for each other in two copies of the node loop): entering an instance swaps the slab ray for the
// object-space one and parks the world one under a marker, popping the marker restores it. The visit order is the
// reference's (bvh::v2: near child first, far child pushed; TLAS leaf -> primitives in order -> each BLAS to
// exhaustion). Measured against the nested form on B200: C2 -5 %, room -14 %, city-10M megakernel -19 %.
// SLAB: the world-space slab constants live in `world_slab` (six words, `slab_stride` apart: a column of shared memory), written
// once per ray; otherwise they are parked on the stack at every instance entry (7 local stores in, 6 local loads out).
template <bool CULL, bool COUNT, bool SLAB = false>
__device__ __forceinline__ void trace_closest(const DScene &sc, V3 wo, V3 wd, Hit &best, uint32_t *stack,
                                              TraceCounters &cnt, float *world_slab = nullptr, int slab_stride = 1)
{
    best.t = SPK_FLT_MAX;
    best.u = 0.f;
    best.v = 0.f;
    best.inst = SPK_INVALID_ID;
    best.face = SPK_INVALID_ID;
    if (sc.num_instances == 0)
    {
        return;
    }
    float slack_abs = 0.f;
    if (CULL)
    {
        slack_abs = sc.cull_extent * 1.52587890625e-5f / fabsf(wd.x); // +inf when d.x == 0: never culls
    }
    const float4 *__restrict__ pairs = sc.pairs;
    const float4 *__restrict__ tris = sc.tris;
    SlabRay sr = make_slab_ray(wo, wd);
    if (SLAB)
    {
        world_slab[0] = sr.ix;
        world_slab[slab_stride] = sr.iy;
        world_slab[2 * slab_stride] = sr.iz;
        world_slab[3 * slab_stride] = sr.ox;
        world_slab[4 * slab_stride] = sr.oy;
        world_slab[5 * slab_stride] = sr.oz;
    }
    V3 lo = wo, ld = wd;             // the ray in the space of the tree being traversed
    uint32_t inst = SPK_INVALID_ID;  // the instance whose BLAS is being traversed; none: TLAS level
    // Both markers have a non-zero low nibble, so the inner loop leaves on them as it does on a leaf and a pop is a
    // bare load: the loop over inner nodes (pops included) is the one piece of code all lanes of a warp stay in.
    stack[0] = kTraversalDone;
    int sp = 1;
    uint32_t top = sc.tlas_root;
    for (;;)
    {
        while ((top & 15u) == 0u)
        {
            const float4 *p = pairs + size_t(top >> 4) * 4;
            float4 l0 = __ldg(p), l1 = __ldg(p + 1), r0 = __ldg(p + 2), r1 = __ldg(p + 3);
            if (COUNT)
            {
                cnt.node_visits++;
            }
            bool hl, hr;
            float tl = slab(sr, l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, hl);
            float tr = slab(sr, r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, hr);
            if (CULL)
            {
                float bound = best.t + (best.t * 2.44140625e-4f + slack_abs);
                hl = hl && tl <= bound;
                hr = hr && tr <= bound;
            }
            uint32_t li = __float_as_uint(l1.z), ri = __float_as_uint(r1.z);
            if (hl)
            {
                uint32_t near_i = li;
                if (hr)
                {
                    uint32_t far_i = ri;
                    if (tl > tr)
                    {
                        near_i = ri;
                        far_i = li;
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
                top = stack[--sp];
            }
        }
        if (top >= kTraversalDone)
        {
            if (top == kTraversalDone)
            {
                return;
            }
            // kLeaveInstance: back at the TLAS level, the world slab constants were parked under the marker
            inst = SPK_INVALID_ID;
            if (SLAB)
            {
                sr.ix = world_slab[0];
                sr.iy = world_slab[slab_stride];
                sr.iz = world_slab[2 * slab_stride];
                sr.ox = world_slab[3 * slab_stride];
                sr.oy = world_slab[4 * slab_stride];
                sr.oz = world_slab[5 * slab_stride];
            }
            else
            {
                sp -= 6;
                sr.ix = __uint_as_float(stack[sp]);
                sr.iy = __uint_as_float(stack[sp + 1]);
                sr.iz = __uint_as_float(stack[sp + 2]);
                sr.ox = __uint_as_float(stack[sp + 3]);
                sr.oy = __uint_as_float(stack[sp + 4]);
                sr.oz = __uint_as_float(stack[sp + 5]);
            }
            sr.nx = signbit(wd.x);
            sr.ny = signbit(wd.y);
            sr.nz = signbit(wd.z);
        }
        else if (inst != SPK_INVALID_ID)
        {
            // BLAS leaf: its triangles are one contiguous run
            const float *fwd = sc.instances[inst].fwd; // row 0 is read when a triangle is hit
            for (uint32_t k = top >> 4, end = (top >> 4) + (top & 15u); k < end; k++)
            {
                if (COUNT)
                {
                    cnt.tri_tests++;
                }
                intersect_triangle(tris + size_t(k) * 3, lo, ld, wo, wd, fwd, inst, best);
            }
        }
        else
        {
            // TLAS leaf: the first primitive now, the rest of the leaf after it (popped right after its BLAS is done)
            const uint32_t i = top >> 4, count = top & 15u;
            if (count > 1u)
            {
                stack[sp++] = ((i + 1u) << 4) | (count - 1u);
            }
            const DInstance *__restrict__ in = sc.instances + i;
            if (__ldg(&in->kind) == 1u)
            {
                if (COUNT)
                {
                    cnt.sphere_tests++;
                }
                intersect_sphere(in, wo, wd, i, best);
            }
            else
            {
                if (COUNT)
                {
                    cnt.inst_visits++;
                }
                // Ray::TransformedBy(inverse): direction is NOT renormalised (Ray.h:23-26)
                float inv[12];
                const float4 *m = reinterpret_cast<const float4 *>(in);
#pragma unroll
                for (int k = 0; k < 3; k++)
                {
                    float4 r = __ldg(m + k);
                    inv[4 * k] = r.x;
                    inv[4 * k + 1] = r.y;
                    inv[4 * k + 2] = r.z;
                    inv[4 * k + 3] = r.w;
                }
                if (SLAB)
                {
                    stack[sp++] = kLeaveInstance;
                }
                else
                {
                    stack[sp] = __float_as_uint(sr.ix);
                    stack[sp + 1] = __float_as_uint(sr.iy);
                    stack[sp + 2] = __float_as_uint(sr.iz);
                    stack[sp + 3] = __float_as_uint(sr.ox);
                    stack[sp + 4] = __float_as_uint(sr.oy);
                    stack[sp + 5] = __float_as_uint(sr.oz);
                    stack[sp + 6] = kLeaveInstance;
                    sp += 7;
                }
                lo = xform_point(inv, wo);
                ld = xform_dir(inv, wd);
                sr = make_slab_ray(lo, ld);
                inst = i;
                top = __ldg(&in->root);
                continue;
            }
        }
        top = stack[--sp];
    }
}

} // namespace spk
