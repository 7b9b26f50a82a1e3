/*
 * spk_traverse_warp.cuh — persistent, warp-scheduled form of the traversal in spk_traverse.cuh for ray batches and
 * ray queues (spk_trace.cu, spk_wavefront.cu).
 *
 * Same arithmetic, same visit order per ray (so the same closest hit, bit for bit); what changes is how the 32 lanes of
 * a warp are kept on the same instruction. spk_traverse.cuh lets every lane run its own nested while loops; on
 * incoherent rays ncu showed 3.5 of 32 lanes active per issued instruction (profiles/r1_trace_thread_per_ray_ncu.txt).
 * Here each lane is a small state machine {top, sp, inst} in one of three states, and every iteration the warp runs the
 * inner step unless enough lanes (VOTE) have accumulated in one of the two leaf states, which then runs instead:
 *   inner   one bvh::v2 inner step: fetch the 64-byte child pair, two slab tests, descend / push / pop
 *   tri     ONE triangle of the BLAS leaf in `top` (a leaf of n triangles is n steps, so short leaves do not wait for
 *           long ones), then the rest of the leaf or pop
 *   tlas    one primitive of the TLAS leaf in `top`: sphere test, or ray into object space and the BLAS root
 * Leaving an instance (the sentinel on the stack) is folded into pop: the world-space slab constants are kept in shared
 * memory per lane, so it costs six shared loads instead of three IEEE divisions in a step of its own.
 * Lanes that finish a ray take the next one from a global counter (persistent threads), so a warp never drains while
 * the batch has rays left.
 * History (ncu, city scene, profiles/r1_wf_extend_city_ncu.txt): the first version voted only between "inner" and "any
 * leaf" and ran the three leaf kinds one after the other with 4.7-6.3 lanes each; safe_inverse alone was 11.5 % of the
 * issued instructions at 4.7 lanes.
 */
#pragma once
#include "spk_traverse.cuh"

namespace spk
{
constexpr uint32_t kSentinel = 0xFFFFFFFFu; // stack marker: "leave the instance, back to the TLAS"
constexpr int kWarpStack = 68;              // 32 TLAS + 1 leaf remainder + 1 sentinel + 32 BLAS (+2 spare)
#ifndef SPK_SMEM_STACK
#define SPK_SMEM_STACK 24
#endif
constexpr int kSmemStack = SPK_SMEM_STACK;              // entries per lane kept in shared memory; deeper ones spill to local memory
constexpr int kLaneSmemWords = kSmemStack + 6; // + the world-space slab constants

// Per-lane traversal stack. The first kSmemStack entries live in shared memory as [depth][lane]: bank = lane whatever
// the depths are, so every push or pop is one conflict-free wavefront (a local-memory stack touches up to 32 lines
// when 32 lanes are at 32 depths, and shares L1 with the node fetches).
struct LaneStack
{
    uint32_t *sm; // this lane's column: sm[depth * 32]; rows kSmemStack.. hold the world slab constants
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
    uint32_t top;   // current node word (inner / leaf)
    int sp;
    uint32_t inst;  // instance being traversed, SPK_INVALID_ID at TLAS level
    float f0, f1, f2, f3; // row 0 of that instance's forward matrix (world_t needs only p.x, Ray.h:43-47)
};

template <bool CULL>
__device__ __forceinline__ void lane_begin(const DScene &sc, LaneState &s, LaneStack &stack, V3 wo, V3 wd)
{
    s.wo = wo;
    s.wd = wd;
    s.lo = wo;
    s.ld = wd;
    s.sr = make_slab_ray(wo, wd);
    float *w = reinterpret_cast<float *>(stack.sm) + kSmemStack * 32;
    w[0] = s.sr.ix;
    w[32] = s.sr.iy;
    w[64] = s.sr.iz;
    w[96] = s.sr.ox;
    w[128] = s.sr.oy;
    w[160] = s.sr.oz;
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

// pop the next item; returns false when the stack is empty (ray finished). A sentinel means BLAS::IntersectInternal
// returned: back to the world ray (its slab constants come from shared memory), then pop again.
__device__ __forceinline__ bool lane_pop(LaneState &s, LaneStack &stack)
{
    for (;;)
    {
        if (s.sp == 0)
        {
            return false;
        }
        s.top = stack.pop(s.sp);
        if (s.top != kSentinel)
        {
            return true;
        }
        s.inst = SPK_INVALID_ID;
        s.lo = s.wo;
        s.ld = s.wd;
        const float *w = reinterpret_cast<const float *>(stack.sm) + kSmemStack * 32;
        s.sr.ix = w[0];
        s.sr.iy = w[32];
        s.sr.iz = w[64];
        s.sr.ox = w[96];
        s.sr.oy = w[128];
        s.sr.oz = w[160];
        s.sr.nx = signbit(s.wd.x);
        s.sr.ny = signbit(s.wd.y);
        s.sr.nz = signbit(s.wd.z);
    }
}

// The top of the hottest tree staged in shared memory (north_star: "the top levels staged in shared memory"): `count` pairs
// starting at pair slot `base`, 64 bytes each. The four 16-byte quarters of pair p are stored rotated by (p >> 1) & 3, so
// that eight lanes reading the same quarter of eight different pairs touch eight different bank groups.
struct TopOfTree
{
    const float4 *sm;
    uint32_t base, count;
};

__device__ __forceinline__ void stage_top_of_tree(const float4 *__restrict__ pairs, float4 *sm, uint32_t base, uint32_t count)
{
    for (uint32_t k = threadIdx.x; k < count * 4u; k += blockDim.x)
    {
        const uint32_t p = k >> 2, q = k & 3u;
        sm[p * 4u + (q ^ ((p >> 1) & 3u))] = __ldg(pairs + size_t(base + p) * 4 + q);
    }
    __syncthreads();
}

// one bvh::v2 inner step (the loop body of trace_closest); returns false when the ray is finished
template <bool CULL, bool COUNT, bool TOP = false>
__device__ __forceinline__ bool lane_inner_step(const float4 *__restrict__ pairs, LaneState &s, LaneStack &stack,
                                                TraceCounters &cnt, const TopOfTree &top = TopOfTree{nullptr, 0, 0})
{
    const uint32_t slot = s.top >> 4;
    float4 l0, l1, r0, r1;
    const uint32_t rel = slot - top.base;
    if (TOP && rel < top.count)
    {
        const float4 *q = top.sm + rel * 4u;
        const uint32_t rot = (rel >> 1) & 3u;
        l0 = q[0u ^ rot];
        l1 = q[1u ^ rot];
        r0 = q[2u ^ rot];
        r1 = q[3u ^ rot];
    }
    else
    {
        const float4 *p = pairs + size_t(slot) * 4;
        ldg256(p, l0, l1);
        ldg256(p + 2, r0, r1);
    }
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

// BLAS leaf in `top`: IntersectTriangle on prim_ids[first], then the rest of the leaf (same order as the leaf loop)
template <bool COUNT>
__device__ __forceinline__ bool lane_tri_step(const DScene &sc, LaneState &s, LaneStack &stack, TraceCounters &cnt)
{
    const uint32_t first = s.top >> 4, count = s.top & 15u;
    const float fwd_row0[4] = {s.f0, s.f1, s.f2, s.f3};
    if (COUNT)
    {
        cnt.tri_tests++;
    }
    intersect_triangle(sc.tris + size_t(first) * 3, s.lo, s.ld, s.wo, s.wd, fwd_row0, s.inst, s.best);
    if (count > 1u)
    {
        s.top = ((first + 1u) << 4) | (count - 1u);
        return true;
    }
    return lane_pop(s, stack);
}

// TLAS leaf in `top`: primitives_[first] now, the rest of the leaf afterwards (same order as the reference's leaf loop)
template <bool COUNT>
__device__ __forceinline__ bool lane_tlas_step(const DScene &sc, LaneState &s, LaneStack &stack, TraceCounters &cnt)
{
    const uint32_t first = s.top >> 4, count = s.top & 15u;
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

/*
 * The persistent loop of one warp. `cursor` is the global ray counter, n the number of rays; fetch(id, o, d, cookie)
 * loads ray id, sets the lane's cookie (the id, or whatever names the ray's output) and returns false to skip the ray,
 * store(cookie, best) writes its closest hit. REFILL: idle lanes wait until this many are
 * idle (or nothing else is left to do) before one lane fetches a block of ray ids with a single atomic.
 */
template <bool CULL, bool COUNT, int REFILL, int VOTE, bool TOP = false, typename Cookie, typename Fetch, typename Store>
__device__ __forceinline__ void warp_trace_loop(const DScene &sc, unsigned long long n, unsigned long long *cursor,
                                                uint32_t *s_lane_words, TraceCounters &cnt, Cookie no_ray, Fetch &&fetch,
                                                Store &&store, const TopOfTree &top = TopOfTree{nullptr, 0, 0})
{
    const unsigned lane = threadIdx.x & 31u;
    const unsigned lt_mask = (1u << lane) - 1u;
    LaneStack stack;
    stack.sm = s_lane_words + (threadIdx.x >> 5) * (kLaneSmemWords * 32) + lane;
    LaneState s;
    s.top = 0u;
    s.inst = SPK_INVALID_ID;
    Cookie cookie = no_ray; // what store() needs to find the ray's output again; no_ray while the lane has none
    bool active = false;
    bool exhausted = false; // warp-uniform: the global counter has passed n
    const float4 *__restrict__ pairs = sc.pairs;
    for (;;)
    {
        // ---- refill idle lanes from the global ray counter
        unsigned idle = __ballot_sync(0xffffffffu, !active);
        if (idle == 0xffffffffu && exhausted)
        {
            break;
        }
        if (!exhausted && (__popc(idle) >= REFILL || idle == 0xffffffffu))
        {
            unsigned long long base = 0;
            const int leader = __ffs(idle) - 1;
            if ((int)lane == leader)
            {
                base = atomicAdd(cursor, (unsigned long long)__popc(idle));
            }
            base = __shfl_sync(0xffffffffu, base, leader);
            if (!active)
            {
                unsigned long long id = base + __popc(idle & lt_mask);
                V3 o, d;
                if (id < n && fetch(id, o, d, cookie))
                {
                    lane_begin<CULL>(sc, s, stack, o, d);
                    active = true;
                }
            }
            exhausted = base + __popc(idle) >= n;
            idle = __ballot_sync(0xffffffffu, !active);
        }
        // ---- majority vote: the state most lanes are in runs one step
        const bool is_inner = active && (s.top & 15u) == 0u;
        const bool is_tri = active && !is_inner && s.inst != SPK_INVALID_ID;
        const bool is_tlas = active && !is_inner && s.inst == SPK_INVALID_ID;
        const unsigned m_tri = __ballot_sync(0xffffffffu, is_tri), m_tlas = __ballot_sync(0xffffffffu, is_tlas);
        const bool no_inner = (~idle & ~(m_tri | m_tlas)) == 0u; // every active lane waits in a leaf state
        const int c_tri = __popc(m_tri), c_tlas = __popc(m_tlas);
        // a leaf state runs as soon as VOTE lanes wait in it (or nothing else can run); inner steps otherwise. A pure
        // majority vote starves the rare state: on soup-1M 6 % of the steps are triangles, their lanes waited until they
        // outnumbered the inner lanes, and inner steps ran with 13.7 lanes (profiles/r1_trace_majority_vote_ncu.txt)
        const bool run_tri = c_tri >= VOTE || (no_inner && c_tri > 0 && c_tri >= c_tlas);
        const bool run_tlas = !run_tri && (c_tlas >= VOTE || (no_inner && c_tlas > 0));
        if (run_tri)
        {
            if (is_tri)
            {
                active = lane_tri_step<COUNT>(sc, s, stack, cnt);
            }
        }
        else if (run_tlas)
        {
            if (is_tlas)
            {
                active = lane_tlas_step<COUNT>(sc, s, stack, cnt);
            }
        }
        else if (is_inner)
        {
            active = lane_inner_step<CULL, COUNT, TOP>(pairs, s, stack, cnt, top);
        }
        // lanes whose stack ran empty: the ray is finished
        if (!active && cookie != no_ray)
        {
            store(cookie, s.best);
            cookie = no_ray;
        }
    }
}
} // namespace spk
