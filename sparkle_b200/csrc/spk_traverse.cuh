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

__device__ __forceinline__ SlabRay make_slab_ray(V3 o, V3 d)
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

// one bvh::v2 traversal over the pair array; LeafFn(first, count) handles a leaf
template <bool CULL, bool COUNT, typename LeafFn>
__device__ __forceinline__ void traverse_tree(const float4 *__restrict__ pairs, uint32_t root, const SlabRay &sr,
                                              uint32_t *stack, const float &best_t, float slack_abs,
                                              TraceCounters &cnt, LeafFn &&leaf)
{
    int sp = 0;
    uint32_t top = root;
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
                float bound = best_t + (best_t * 2.44140625e-4f + slack_abs);
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
                goto pop;
            }
        }
        leaf(top >> 4, top & 15u);
    pop:
        if (sp == 0)
        {
            return;
        }
        top = stack[--sp];
    }
}

// SceneRenderProxy::Intersect<false> up to the winning candidate (GetIntersection is done by the shading stage)
template <bool CULL, bool COUNT>
__device__ __forceinline__ void trace_closest(const DScene &sc, V3 wo, V3 wd, Hit &best, uint32_t *stack,
                                              TraceCounters &cnt)
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
    const SlabRay wsr = make_slab_ray(wo, wd);
    const float4 *__restrict__ pairs = sc.pairs;
    traverse_tree<CULL, COUNT>(pairs, sc.tlas_root, wsr, stack, best.t, slack_abs, cnt, [&](uint32_t first,
                                                                                              uint32_t count) {
        for (uint32_t i = first; i < first + count; i++)
        {
            const DInstance *__restrict__ in = sc.instances + i;
            if (__ldg(&in->kind) == 1u)
            {
                if (COUNT)
                {
                    cnt.sphere_tests++;
                }
                intersect_sphere(in, wo, wd, i, best);
                continue;
            }
            if (COUNT)
            {
                cnt.inst_visits++;
            }
            // Ray::TransformedBy(inverse): direction is NOT renormalised (Ray.h:23-26)
            float inv[12], fwd[12];
            const float4 *m = reinterpret_cast<const float4 *>(in);
#pragma unroll
            for (int k = 0; k < 3; k++)
            {
                float4 a = __ldg(m + k);
                inv[4 * k] = a.x;
                inv[4 * k + 1] = a.y;
                inv[4 * k + 2] = a.z;
                inv[4 * k + 3] = a.w;
                float4 b = __ldg(m + 3 + k);
                fwd[4 * k] = b.x;
                fwd[4 * k + 1] = b.y;
                fwd[4 * k + 2] = b.z;
                fwd[4 * k + 3] = b.w;
            }
            V3 lo = xform_point(inv, wo);
            V3 ld = xform_dir(inv, wd);
            const SlabRay lsr = make_slab_ray(lo, ld);
            const float4 *__restrict__ tris = sc.tris;
            traverse_tree<CULL, COUNT>(pairs, __ldg(&in->root), lsr, stack + 32, best.t, slack_abs, cnt,
                                       [&](uint32_t tfirst, uint32_t tcount) {
                                           for (uint32_t k = tfirst; k < tfirst + tcount; k++)
                                           {
                                               if (COUNT)
                                               {
                                                   cnt.tri_tests++;
                                               }
                                               intersect_triangle(tris + size_t(k) * 3, lo, ld, wo, wd, fwd, i, best);
                                           }
                                       });
        }
    });
}
} // namespace spk
