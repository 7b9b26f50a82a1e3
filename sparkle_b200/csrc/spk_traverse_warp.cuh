/*
 * spk_traverse_warp.cuh — warp-converged form of the traversal in spk_traverse.cuh for incoherent ray batches.
 *
 * Same arithmetic, same visit order per ray (so the same closest hit, bit for bit); what changes is how the 32 lanes of
 * a warp are kept on the same instruction. spk_traverse.cuh lets every lane run its own nested while loops; on
 * incoherent rays ncu showed 3.5 of 32 lanes active per issued instruction (profiles/r1_trace_simple.txt). Here each lane
 * is a small state machine {top, sp, level} and the warp alternates between two phases it votes on:
 *   inner phase  every lane whose `top` is an inner node does one step (fetch the 64-byte child pair, two slab tests,
 *                descend / push / pop); repeated while at least one lane has inner work and fewer than kLeafVote
 *                lanes are parked on a leaf;
 *   leaf phase   every lane parked on a leaf item handles it: BLAS leaf -> its triangles, TLAS leaf -> enter the next
 *                instance (sphere test, or ray into object space and BLAS root), sentinel -> back to world space.
 * Lanes that finish a ray take the next one from a global counter (persistent threads), so a warp never drains while
 * the batch has rays left.
 */
#pragma once
#include "spk_traverse.cuh"

namespace spk
{
constexpr uint32_t kSentinel = 0xFFFFFFFFu; // stack marker: "leave the instance, back to the TLAS"
constexpr int kWarpStack = 68;              // 32 TLAS + 1 leaf remainder + 1 sentinel + 32 BLAS (+2 spare)
constexpr int kSmemStack = 24;              // entries per lane kept in shared memory; deeper ones spill to local memory

// Per-lane traversal stack. A stack in local memory shares L1 with the node fetches, and 32 lanes at 32 different
// depths touch 32 different lines per push or pop (ncu: L1/TEX at 86 % of peak). The first kSmemStack
// entries therefore live in shared memory as [depth][lane]: bank = lane whatever the depths are, so every push or pop
// is one conflict-free wavefront.
struct LaneStack
{
    uint32_t *sm; // this lane's column: sm[depth * 32]
    uint32_t deep[kWarpStack - kSmemStack];
    __device__ __forceinline__ void push(int &sp, uint32_t v)
    {
        if (sp < kSmemStack)
        {
            sm[sp * 32] = v;
        }
        else
        {
            deep[sp - kSmemStack] = v;
        }
        sp++;
    }
    __device__ __forceinline__ uint32_t pop(int &sp)
    {
        --sp;
        return sp < kSmemStack ? sm[sp * 32] : deep[sp - kSmemStack];
    }
};

struct LaneState
{
    V3 wo, wd;      // world ray
    V3 lo, ld;      // ray of the current level (object space inside an instance)
    SlabRay sr;     // slab constants of the current level
    Hit best;
    float slack_abs;
    uint32_t top;   // current node word (inner / leaf) or kSentinel
    int sp;
    uint32_t inst;  // instance being traversed, SPK_INVALID_ID at TLAS level
    float f0, f1, f2, f3; // row 0 of that instance's forward matrix (world_t needs only p.x, Ray.h:43-47)
};

template <bool CULL> __device__ __forceinline__ void lane_begin(const DScene &sc, LaneState &s, V3 wo, V3 wd)
{
    s.wo = wo;
    s.wd = wd;
    s.lo = wo;
    s.ld = wd;
    s.sr = make_slab_ray(wo, wd);
    s.best.t = SPK_FLT_MAX;
    s.best.u = 0.f;
    s.best.v = 0.f;
    s.best.inst = SPK_INVALID_ID;
    s.best.face = SPK_INVALID_ID;
    s.slack_abs = CULL ? sc.cull_extent * 1.52587890625e-5f / fabsf(wd.x) : 0.f;
    s.top = sc.tlas_root;
    s.sp = 0;
    s.inst = SPK_INVALID_ID;
}

// pop the next item; returns false when the stack is empty (ray finished)
__device__ __forceinline__ bool lane_pop(LaneState &s, LaneStack &stack)
{
    if (s.sp == 0)
    {
        return false;
    }
    s.top = stack.pop(s.sp);
    return true;
}

// one bvh::v2 inner step (identical to the loop body of traverse_tree); returns false when the ray is finished
template <bool CULL, bool COUNT>
__device__ __forceinline__ bool lane_inner_step(const float4 *__restrict__ pairs, LaneState &s, LaneStack &stack,
                                                TraceCounters &cnt)
{
    const float4 *p = pairs + size_t(s.top >> 4) * 4;
    float4 l0 = __ldg(p), l1 = __ldg(p + 1), r0 = __ldg(p + 2), r1 = __ldg(p + 3);
    if (COUNT)
    {
        cnt.node_visits++;
    }
    bool hl, hr;
    float tl = slab(s.sr, l0.x, l0.y, l0.z, l0.w, l1.x, l1.y, hl);
    float tr = slab(s.sr, r0.x, r0.y, r0.z, r0.w, r1.x, r1.y, hr);
    if (CULL)
    {
        float bound = s.best.t + (s.best.t * 2.44140625e-4f + s.slack_abs);
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
            stack.push(s.sp, far_i);
        }
        s.top = near_i;
        return true;
    }
    if (hr)
    {
        s.top = ri;
        return true;
    }
    return lane_pop(s, stack);
}

// handle the leaf item in `top`; returns false when the ray is finished
template <bool CULL, bool COUNT>
__device__ __forceinline__ bool lane_leaf_step(const DScene &sc, LaneState &s, LaneStack &stack, TraceCounters &cnt)
{
    if (s.top == kSentinel)
    {
        // BLAS::IntersectInternal returned: back to the world ray
        s.inst = SPK_INVALID_ID;
        s.lo = s.wo;
        s.ld = s.wd;
        s.sr = make_slab_ray(s.wo, s.wd);
        return lane_pop(s, stack);
    }
    const uint32_t first = s.top >> 4, count = s.top & 15u;
    if (s.inst != SPK_INVALID_ID)
    {
        // BLAS leaf: IntersectTriangle over prim_ids[first, first+count)
        const float fwd_row0[4] = {s.f0, s.f1, s.f2, s.f3};
        const float4 *__restrict__ tris = sc.tris;
        for (uint32_t k = first; k < first + count; k++)
        {
            if (COUNT)
            {
                cnt.tri_tests++;
            }
            intersect_triangle(tris + size_t(k) * 3, s.lo, s.ld, s.wo, s.wd, fwd_row0, s.inst, s.best);
        }
        return lane_pop(s, stack);
    }
    // TLAS leaf: primitives_[first] now, the rest of the leaf afterwards (same order as the reference's leaf loop)
    if (count > 1u)
    {
        stack.push(s.sp, ((first + 1u) << 4) | (count - 1u));
    }
    const DInstance *__restrict__ in = sc.instances + first;
    if (__ldg(&in->kind) == 1u)
    {
        if (COUNT)
        {
            cnt.sphere_tests++;
        }
        intersect_sphere(in, s.wo, s.wd, first, s.best);
        return lane_pop(s, stack);
    }
    if (COUNT)
    {
        cnt.inst_visits++;
    }
    float inv[12];
    const float4 *m = reinterpret_cast<const float4 *>(in);
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        float4 a = __ldg(m + k);
        inv[4 * k] = a.x;
        inv[4 * k + 1] = a.y;
        inv[4 * k + 2] = a.z;
        inv[4 * k + 3] = a.w;
    }
    float4 f = __ldg(m + 3);
    s.f0 = f.x;
    s.f1 = f.y;
    s.f2 = f.z;
    s.f3 = f.w;
    s.lo = xform_point(inv, s.wo); // Ray::TransformedBy(inverse), direction not renormalised
    s.ld = xform_dir(inv, s.wd);
    s.sr = make_slab_ray(s.lo, s.ld);
    s.inst = first;
    stack.push(s.sp, kSentinel);
    s.top = __ldg(&in->root);
    return true;
}
} // namespace spk
