/*
 * spk_wavefront.cu — wavefront (queue) form of the path tracer, selected by SPK_FLAG_WAVEFRONT.
 *
 * Same per-sample arithmetic as render_kernel (spk_render.cu; CPURenderer.cpp:220-450): every (pixel, sample) path draws
 * from the same seeded stream and runs the same device functions, so each sample's radiance is bit-identical to the
 * megakernel's; only the ORDER in which samples are added to a pixel differs (vector float atomics), i.e. float
 * association of the sum (~1e-7 relative). SPK_ACCUM_SUM only.
 *
 * A pool of P path slots lives in HBM (one 96-byte record per slot, see WfState). One iteration = four launches, no host
 * round trip (counts live in a device control block):
 *   wf_shade    one thread per traced slot. The CTA first sorts its 256 slots by material class in shared memory
 *               (miss / sphere|mesh x pbr|dielectric x textured x emissive), so a warp shades like with like.
 *               Continuing paths are appended to the next ray queue, finished ones to the regen list
 *               (warp-ballot compaction, one atomic per warp).
 *   wf_regen    finished slots take the next (pixel, sample) work items in order: consecutive threads get consecutive
 *               pixels of a row, so new primary rays enter the queue as coherent runs.
 *   wf_advance  one thread: rotates the counters.
 *   wf_extend   persistent warp-voting traversal (spk_traverse_warp.cuh) over the ray queue: lanes refill from the
 *               queue, so warps stay full however long individual rays take.
 */
#include "spk_context.h"

#include <cstdlib>
#include <cub/device/device_radix_sort.cuh>
#include "spk_shade.cuh"
#include "spk_traverse.cuh"
#include "spk_traverse_warp.cuh"

namespace spk
{
namespace
{
constexpr int kShadeThreads = 256;
constexpr int kExtendWarps = 4;
constexpr uint32_t kEmptySlot = 0xFFFFFFFFu;
constexpr size_t kRec = 6; // float4 per slot record
// The slot records stream through the memory system once per iteration (1.6 GB for 16 Mi slots): they are read and
// written with the evict-first hint so that they do not push the BVH and the mesh attributes out of the L2.
#ifndef SPK_WF_STREAM
#define SPK_WF_STREAM 1
#endif
#if SPK_WF_STREAM
#define SPK_LD(p) __ldcs(p)
#define SPK_ST(p, v) __stcs(p, v)
#else
#define SPK_LD(p) (*(p))
#define SPK_ST(p, v) (*(p) = (v))
#endif

struct WfCtrl
{
    uint32_t n_queue[2]; // entries in queue[0] / queue[1]
    uint32_t n_regen;    // entries in the regen list
    uint32_t pad;
    unsigned long long next_ray; // wf_extend's queue cursor
    unsigned long long work_next, work_total; // (pixel, sample) items handed out so far / in this call
    unsigned long long rays;                   // segments traced in this call
};

struct WfState
{
    // One 96-byte record per slot, three 32-byte sectors that are always read and written whole (the slot order of a queue
    // is close to random after a few bounces: with one array per field every 4- or 16-byte access cost a sector of its own
    // and partial sector writes cost a fill on top, 586 B of DRAM traffic per shaded slot):
    //   [0] {origin.xyz, pixel bits} [1] {direction.xyz, bounce bits}      written by regen / shade, read by extend / shade
    //   [2] {t, u, v, instance slot bits} [3] {face bits, -, -, -}         written by extend, read by shade
    //   [4] rng state [5] {throughput.xyz, -}                              written by regen / shade, read by shade
    float4 *rec;
    uint32_t *queue[2];
    uint32_t *regen;
    uint32_t *keys;  // sort key of every entry of the queue being filled (ray sorting on; else unused)
    WfCtrl *ctrl;
    uint32_t slots;
};

struct WfArgs
{
    uint32_t width, height;
    uint32_t first_sample, sample_stride, max_bounce;
    float output_limit;
    // ray sorting (SPK_OPT_WAVEFRONT_SORT): cells of the scene box, key_bits per axis; 0 = off
    uint32_t key_bits;
    float key_lo[3], key_scale[3];
};

// Sort key of a queued ray: Morton code of the origin's cell in the scene box, then the direction's octant. Rays that start
// in the same cell and head into the same octant walk the same instances and the same side of their trees.
__device__ __forceinline__ uint32_t spread3(uint32_t v) // 10 bits -> every third bit
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}
__device__ __forceinline__ uint32_t ray_sort_key(const WfArgs &a, V3 o, V3 d)
{
    const float top = (float)((1u << a.key_bits) - 1u);
    const uint32_t cx = (uint32_t)fminf(fmaxf((o.x - a.key_lo[0]) * a.key_scale[0], 0.f), top);
    const uint32_t cy = (uint32_t)fminf(fmaxf((o.y - a.key_lo[1]) * a.key_scale[1], 0.f), top);
    const uint32_t cz = (uint32_t)fminf(fmaxf((o.z - a.key_lo[2]) * a.key_scale[2], 0.f), top);
    const uint32_t cell = (spread3(cx) << 2) | (spread3(cy) << 1) | spread3(cz);
    const uint32_t oct = (d.x < 0.f ? 4u : 0u) | (d.y < 0.f ? 2u : 0u) | (d.z < 0.f ? 1u : 0u);
    return (cell << 3) | oct;
}

__global__ void wf_init_kernel(WfState st, unsigned long long total)
{
    uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k < st.slots)
    {
        st.regen[k] = k;
    }
    if (k == 0)
    {
        WfCtrl c;
        c.n_queue[0] = c.n_queue[1] = 0;
        c.n_regen = st.slots;
        c.next_ray = 0;
        c.work_next = 0;
        c.work_total = total;
        c.rays = 0;
        *st.ctrl = c;
    }
}

// RenderPixel + SetupViewRay (CPURenderer.cpp:398-417, 220-238) for work item w = sample * W*H + pixel
__global__ void __launch_bounds__(kShadeThreads)
    wf_regen_kernel(const DCamera cam, const WfArgs a, WfState st, uint32_t nxt)
{
    const WfCtrl c = *st.ctrl;
    const unsigned long long left = c.work_total - c.work_next;
    const uint32_t gen = uint32_t(c.n_regen < left ? c.n_regen : left); // slots that get a new path
    const uint32_t npix = a.width * a.height;
    const uint32_t s0 = uint32_t(c.work_next / npix), p0 = uint32_t(c.work_next % npix);
    const float pw = 1.f / (float)(a.width - 1), ph = 1.f / (float)(a.height - 1);
    const V3 cpos = mk(cam.position[0], cam.position[1], cam.position[2]);
    const V3 cup = mk(cam.up[0], cam.up[1], cam.up[2]);
    const V3 cright = mk(cam.right[0], cam.right[1], cam.right[2]);
    const V3 max_u = mk(cam.max_u[0], cam.max_u[1], cam.max_u[2]);
    const V3 max_v = mk(cam.max_v[0], cam.max_v[1], cam.max_v[2]);
    const V3 lower_left = mk(cam.lower_left[0], cam.lower_left[1], cam.lower_left[2]);
    // a fixed grid strides over the list (its length is only known on the device)
    const uint32_t stride = gridDim.x * blockDim.x;
    if (a.key_bits)
    {
        // queue positions behind the last entry this iteration will have: largest key, so the sort leaves them at the end
        for (uint32_t k = c.n_queue[nxt] + gen + blockIdx.x * blockDim.x + threadIdx.x; k < st.slots; k += stride)
        {
            st.keys[k] = 0xFFFFFFFFu;
        }
    }
    for (uint32_t k = blockIdx.x * blockDim.x + threadIdx.x; k < c.n_regen; k += stride)
    {
        const uint32_t slot = st.regen[k];
        if (k >= gen)
        {
            continue; // no work left for this slot: it stays out of the queues from here on
        }
        // work item w = work_next + k = sample * W*H + pixel, without a 64-bit division per path
        uint32_t s = s0;
        unsigned long long p64 = (unsigned long long)p0 + k;
        while (p64 >= npix)
        {
            p64 -= npix;
            s++;
        }
        const uint32_t p = uint32_t(p64);
        const uint32_t i = p % a.width, j = p / a.width;
        const uint32_t frame_seed = a.first_sample + s * a.sample_stride; // CPURenderer.cpp:442-443, unsigned wrap
        Rng rng = rng_seed(j * a.width + i + frame_seed * a.width * a.height);
        float u = ((float)i + rng_unit(rng)) * pw;
        float v = ((float)j + rng_unit(rng)) * ph;
        float dx, dy;
        unit_disk(rng, dx, dy);
        dx = dx * cam.aperture_radius;
        dy = dy * cam.aperture_radius;
        V3 origin = cpos + (dx * cright + dy * cup);
        V3 target = lower_left + u * max_u + v * max_v;
        V3 dir = normalized(target - origin);
        float4 *r = st.rec + kRec * size_t(slot);
        SPK_ST(r + 0, make_float4(origin.x, origin.y, origin.z, __uint_as_float(p)));
        SPK_ST(r + 1, make_float4(dir.x, dir.y, dir.z, __uint_as_float(0u)));
        SPK_ST(r + 4, make_float4(__uint_as_float(rng.s0), __uint_as_float(rng.s1), __uint_as_float(rng.s2), __uint_as_float(rng.s3)));
        SPK_ST(r + 5, make_float4(1.f, 1.f, 1.f, 0.f));
        st.queue[nxt][c.n_queue[nxt] + k] = slot; // after the entries wf_shade appended; wf_advance adds the count
        if (a.key_bits)
        {
            st.keys[c.n_queue[nxt] + k] = ray_sort_key(a, origin, dir);
        }
    }
}

__global__ void wf_advance_kernel(WfState st, uint32_t cur, unsigned long long *__restrict__ counters)
{
    WfCtrl *c = st.ctrl;
    const uint32_t nxt = cur ^ 1u;
    unsigned long long left = c->work_total - c->work_next;
    unsigned long long gen = c->n_regen < left ? c->n_regen : left;
    c->n_queue[nxt] += uint32_t(gen);
    c->work_next += gen;
    c->n_regen = 0;
    c->n_queue[cur] = 0;
    c->next_ray = 0;
    c->rays += c->n_queue[nxt];
    counters[0] += gen;               // paths started
    counters[1] += c->n_queue[nxt];   // segments traced
}

// ---- extend: closest hit of every queued ray; the loop of trace_persistent_kernel (spk_trace.cu) over an indirect queue
template <bool CULL, bool COUNT, int REFILL>
#ifndef SPK_WF_EXTEND_CTAS
#define SPK_WF_EXTEND_CTAS 8 // 64 registers instead of 72: 8 CTAs per SM instead of 7 (city-10M 4K x 16 spp: 403 -> 385 ms)
#endif
__global__ void __launch_bounds__(kExtendWarps * 32, SPK_WF_EXTEND_CTAS)
    wf_extend_kernel(const DScene sc, WfState st, uint32_t q, unsigned long long *__restrict__ counters)
{
    __shared__ uint32_t s_lane_words[kExtendWarps * kLaneSmemWords * 32];
    const uint32_t *__restrict__ queue = st.queue[q];
    const float4 *rec = st.rec;
    TraceCounters cnt{0, 0, 0, 0};
#ifndef SPK_WF_VOTE
#define SPK_WF_VOTE 8
#endif
    warp_trace_loop<CULL, COUNT, REFILL, SPK_WF_VOTE>(
        sc, st.ctrl->n_queue[q], &st.ctrl->next_ray, s_lane_words, cnt, kEmptySlot,
        [&](unsigned long long id, V3 &o, V3 &d, uint32_t &slot_out) {
            const uint32_t slot = __ldg(queue + id);
            float4 ro = SPK_LD(rec + kRec * size_t(slot)), rd = SPK_LD(rec + kRec * size_t(slot) + 1);
            o = mk(ro.x, ro.y, ro.z);
            d = mk(rd.x, rd.y, rd.z);
            slot_out = slot; // kept by the lane: the store used to read the queue again, a dependent load per finished ray
            return true;
        },
        [&](uint32_t slot, const Hit &best) {
            float4 *r = st.rec + kRec * size_t(slot) + 2;
            SPK_ST(r + 0, make_float4(best.t, best.u, best.v, __uint_as_float(best.inst)));
            SPK_ST(r + 1, make_float4(__uint_as_float(best.face), 0.f, 0.f, 0.f)); // the whole sector: no fill for a partial write
        });
    if (COUNT)
    {
        const unsigned lane = threadIdx.x & 31u;
        uint32_t v[4] = {cnt.node_visits, cnt.tri_tests, cnt.inst_visits, cnt.sphere_tests};
#pragma unroll
        for (int k = 0; k < 4; k++)
        {
            uint32_t sum = __reduce_add_sync(0xffffffffu, v[k]);
            if (lane == 0 && sum)
            {
                atomicAdd(counters + 2 + k, (unsigned long long)sum);
            }
        }
    }
}

// ---- shade: SamplePixel's loop body after Intersect (CPURenderer.cpp:259-369) for every traced slot
#ifndef SPK_WF_SHADE_CTAS
#define SPK_WF_SHADE_CTAS 3 // 80 registers: three CTAs per SM (the sort key would take it to 86 and two)
#endif
__global__ void __launch_bounds__(kShadeThreads, SPK_WF_SHADE_CTAS)
    wf_shade_kernel(const DScene sc, const WfArgs a, WfState st, uint32_t cur, float4 *__restrict__ accum)
{
    __shared__ uint32_t s_bin[32];
    // the slot state travels through shared memory to the thread that shades it: [word][sorted position]
    constexpr int kStateWords = 22; // slot, ray 8, hit 4, face, throughput 4 (w unused), rng 4
    __shared__ uint32_t s_state[kStateWords * kShadeThreads];
    const uint32_t tid = threadIdx.x, lane = tid & 31u;
    const uint32_t k = blockIdx.x * kShadeThreads + tid;
    const uint32_t n = st.ctrl->n_queue[cur];
    if (blockIdx.x * kShadeThreads >= n)
    {
        return; // whole CTA beyond the queue
    }
    // -- every thread fetches ALL the state of the slot at its queue position in one round of independent loads (the class
    //    used to be fetched first, the rest after the sort and the rng after the emissive test: three dependent trips to
    //    HBM per path), then the CTA counting-sorts the slots by material class and hands the state over in shared memory
    if (tid < 32)
    {
        s_bin[tid] = 0;
    }
    __syncthreads();
    uint32_t my = kEmptySlot, key = 31u, pos = 0;
    float4 q_ro = make_float4(0.f, 0.f, 0.f, 0.f), q_rd = q_ro, q_h = q_ro, q_t = q_ro;
    uint4 q_rng = make_uint4(0u, 0u, 0u, 0u);
    uint32_t q_face = 0;
    if (k < n)
    {
        my = st.queue[cur][k];
        const float4 *r = st.rec + kRec * size_t(my);
        q_ro = SPK_LD(r + 0);
        q_rd = SPK_LD(r + 1);
        q_h = SPK_LD(r + 2);
        q_face = __float_as_uint(SPK_LD(r + 3).x);
        const float4 g = SPK_LD(r + 4);
        q_rng = make_uint4(__float_as_uint(g.x), __float_as_uint(g.y), __float_as_uint(g.z), __float_as_uint(g.w));
        q_t = SPK_LD(r + 5);
        key = shade_class(sc, __float_as_uint(q_h.w));
    }
    pos = atomicAdd(&s_bin[key], 1u);
    __syncthreads();
    if (tid < 32)
    {
        uint32_t c = s_bin[tid], incl = c;
#pragma unroll
        for (int d = 1; d < 32; d <<= 1)
        {
            uint32_t t = __shfl_up_sync(0xffffffffu, incl, d);
            if ((int)lane >= d)
            {
                incl += t;
            }
        }
        s_bin[tid] = incl - c; // exclusive start of the class
    }
    __syncthreads();
    {
        uint32_t *o = s_state + s_bin[key] + pos;
        o[0 * kShadeThreads] = my;
        o[1 * kShadeThreads] = __float_as_uint(q_ro.x);
        o[2 * kShadeThreads] = __float_as_uint(q_ro.y);
        o[3 * kShadeThreads] = __float_as_uint(q_ro.z);
        o[4 * kShadeThreads] = __float_as_uint(q_ro.w);
        o[5 * kShadeThreads] = __float_as_uint(q_rd.x);
        o[6 * kShadeThreads] = __float_as_uint(q_rd.y);
        o[7 * kShadeThreads] = __float_as_uint(q_rd.z);
        o[8 * kShadeThreads] = __float_as_uint(q_rd.w);
        o[9 * kShadeThreads] = __float_as_uint(q_h.x);
        o[10 * kShadeThreads] = __float_as_uint(q_h.y);
        o[11 * kShadeThreads] = __float_as_uint(q_h.z);
        o[12 * kShadeThreads] = __float_as_uint(q_h.w);
        o[13 * kShadeThreads] = q_face;
        o[14 * kShadeThreads] = __float_as_uint(q_t.x);
        o[15 * kShadeThreads] = __float_as_uint(q_t.y);
        o[16 * kShadeThreads] = __float_as_uint(q_t.z);
        o[17 * kShadeThreads] = q_rng.x;
        o[18 * kShadeThreads] = q_rng.y;
        o[19 * kShadeThreads] = q_rng.z;
        o[20 * kShadeThreads] = q_rng.w;
    }
    __syncthreads();
    const uint32_t *in = s_state + tid;
    const uint32_t slot = in[0];

    bool done = false, alive = slot != kEmptySlot;
    uint32_t sort_key = 0;
    V3 color = mk(0.f, 0.f, 0.f);
    float valid = 1.f;
    uint32_t pixel = 0;
    if (alive)
    {
        const float4 ro = make_float4(__uint_as_float(in[1 * kShadeThreads]), __uint_as_float(in[2 * kShadeThreads]),
                                      __uint_as_float(in[3 * kShadeThreads]), __uint_as_float(in[4 * kShadeThreads]));
        const float4 rd = make_float4(__uint_as_float(in[5 * kShadeThreads]), __uint_as_float(in[6 * kShadeThreads]),
                                      __uint_as_float(in[7 * kShadeThreads]), __uint_as_float(in[8 * kShadeThreads]));
        pixel = __float_as_uint(ro.w);
        uint32_t bounce = __float_as_uint(rd.w);
        V3 origin = mk(ro.x, ro.y, ro.z), dir = mk(rd.x, rd.y, rd.z);
        V3 throughput = mk(__uint_as_float(in[14 * kShadeThreads]), __uint_as_float(in[15 * kShadeThreads]),
                           __uint_as_float(in[16 * kShadeThreads]));
        Hit hit;
        hit.t = __uint_as_float(in[9 * kShadeThreads]);
        hit.u = __uint_as_float(in[10 * kShadeThreads]);
        hit.v = __uint_as_float(in[11 * kShadeThreads]);
        hit.inst = in[12 * kShadeThreads];
        if (hit.inst == SPK_INVALID_ID)
        {
            if (sc.sky.has_sky)
            {
                color = cmul(sky_evaluate(sc, dir), throughput);
            }
            done = true;
        }
        else
        {
            hit.face = in[13 * kShadeThreads];
            Surface sf;
            get_intersection(sc, origin, dir, hit, sf);
            V3 emissive = material_emissive(sc, sc.materials[sf.material], sf.u, sf.v);
            if (norm(emissive) > kEps)
            {
                color = cmul(emissive, throughput);
                done = true;
            }
            else
            {
                Rng rng{in[17 * kShadeThreads], in[18 * kShadeThreads], in[19 * kShadeThreads], in[20 * kShadeThreads]};
                V3 next_dir;
                V3 this_throughput = sample_surface(sc, rng, sf, dir, next_dir);
                throughput = cmul(throughput, this_throughput);
                if (sqnorm(throughput) < kEps)
                {
                    valid = -1.f;
                    done = true;
                }
                else
                {
                    if (bounce >= 3)
                    {
                        float p = fmaxf(throughput.x, fmaxf(throughput.y, throughput.z));
                        p = p < 0.05f ? 0.05f : (p > 1.0f ? 1.0f : p); // std::clamp
                        if (rng_unit(rng) > p)
                        {
                            done = true;
                        }
                        else
                        {
                            throughput = throughput / p;
                        }
                    }
                    if (!done)
                    {
                        bounce++;
                        done = bounce >= a.max_bounce; // the bounce cap adds nothing (CPURenderer.cpp:256)
                    }
                    if (!done)
                    {
                        origin = sf.location + next_dir * kTolerance;
                        float4 *r = st.rec + kRec * size_t(slot);
                        SPK_ST(r + 0, make_float4(origin.x, origin.y, origin.z, ro.w));
                        SPK_ST(r + 1, make_float4(next_dir.x, next_dir.y, next_dir.z, __uint_as_float(bounce)));
                        SPK_ST(r + 4, make_float4(__uint_as_float(rng.s0), __uint_as_float(rng.s1), __uint_as_float(rng.s2),
                                                  __uint_as_float(rng.s3)));
                        SPK_ST(r + 5, make_float4(throughput.x, throughput.y, throughput.z, 0.f));
                        if (a.key_bits)
                        {
                            sort_key = ray_sort_key(a, origin, next_dir);
                        }
                    }
                }
            }
        }
    }
    // -- finished paths: clamp and add (RenderPixel, CPURenderer.cpp:411-416); the slot goes to the regen list
    if (alive && done)
    {
        float4 add = make_float4(fminf(color.x, a.output_limit), fminf(color.y, a.output_limit),
                                 fminf(color.z, a.output_limit), valid);
        atomicAdd(accum + pixel, add); // one 16-byte vector reduction (sm_90+)
    }
    const uint32_t nxt = cur ^ 1u;
    const unsigned m_cont = __ballot_sync(0xffffffffu, alive && !done);
    const unsigned m_done = __ballot_sync(0xffffffffu, alive && done);
    const unsigned lt = (1u << lane) - 1u;
    uint32_t base_c = 0, base_d = 0;
    if (lane == 0)
    {
        if (m_cont)
        {
            base_c = atomicAdd(&st.ctrl->n_queue[nxt], (uint32_t)__popc(m_cont));
        }
        if (m_done)
        {
            base_d = atomicAdd(&st.ctrl->n_regen, (uint32_t)__popc(m_done));
        }
    }
    base_c = __shfl_sync(0xffffffffu, base_c, 0);
    base_d = __shfl_sync(0xffffffffu, base_d, 0);
    if (alive && !done)
    {
        st.queue[nxt][base_c + __popc(m_cont & lt)] = slot;
        if (a.key_bits)
        {
            st.keys[base_c + __popc(m_cont & lt)] = sort_key;
        }
    }
    else if (alive)
    {
        st.regen[base_d + __popc(m_done & lt)] = slot;
    }
}
} // namespace

struct WfSort // ray sorting: a third queue buffer, the sorted keys and cub's scratch, behind the slot arrays
{
    uint32_t *queue_spare = nullptr, *keys_sorted = nullptr;
    void *tmp = nullptr;
    size_t tmp_bytes = 0;
    int end_bit = 0;
};

static int wf_reserve(SpkContext *ctx, uint32_t slots, WfState &st, WfSort &sort, uint32_t key_bits)
{
    // one allocation, carved into 256-byte aligned arrays. The default pool is large (64 Mi slots = 7.2 GB): when the device
    // cannot give that much the pool is halved until it fits (fewer slots only mean more, shorter iterations).
    auto al = [](size_t b) { return (b + 255) & ~size_t(255); };
    size_t b_rec = 0, b_u32 = 0;
    sort.end_bit = int(3 * key_bits + 3);
    for (;;)
    {
        b_rec = al(size_t(slots) * kRec * 16);
        b_u32 = al(size_t(slots) * 4);
        size_t b_sort = 0;
        if (key_bits)
        {
            SPK_CUDA_TRY(cub::DeviceRadixSort::SortPairs(nullptr, sort.tmp_bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                                         (const uint32_t *)nullptr, (uint32_t *)nullptr, slots, 0, sort.end_bit,
                                                         ctx->stream));
            b_sort = 3 * b_u32 + al(sort.tmp_bytes);
        }
        const size_t total = b_rec + b_u32 * 3 + 256 + b_sort;
        if (ctx->wf_slots == slots && ctx->wf_bytes >= total)
        {
            break;
        }
        cudaFree(ctx->wf_block);
        ctx->wf_block = nullptr;
        ctx->wf_slots = 0;
        ctx->wf_bytes = 0;
        const cudaError_t e = cudaMalloc(&ctx->wf_block, total);
        if (e == cudaSuccess)
        {
            ctx->wf_slots = slots;
            ctx->wf_bytes = total;
            break;
        }
        cudaGetLastError(); // clear the allocation failure
        ctx->wf_block = nullptr;
        if (e != cudaErrorMemoryAllocation || slots <= (1u << 20))
        {
            set_error(std::string("spk_cuda_render: wavefront slot pool: ") + cudaGetErrorString(e));
            return e == cudaErrorMemoryAllocation ? SPK_ERR_OUT_OF_MEMORY : SPK_ERR_CUDA;
        }
        slots = ((slots / 2u) + 255u) & ~255u;
    }
    char *p = static_cast<char *>(ctx->wf_block);
    st.rec = reinterpret_cast<float4 *>(p);
    p += b_rec;
    st.queue[0] = reinterpret_cast<uint32_t *>(p);
    p += b_u32;
    st.queue[1] = reinterpret_cast<uint32_t *>(p);
    p += b_u32;
    st.regen = reinterpret_cast<uint32_t *>(p);
    p += b_u32;
    st.ctrl = reinterpret_cast<WfCtrl *>(p);
    p += 256;
    st.keys = nullptr;
    if (key_bits)
    {
        st.keys = reinterpret_cast<uint32_t *>(p);
        p += b_u32;
        sort.keys_sorted = reinterpret_cast<uint32_t *>(p);
        p += b_u32;
        sort.queue_spare = reinterpret_cast<uint32_t *>(p);
        p += b_u32;
        sort.tmp = p;
    }
    st.slots = slots;
    return SPK_OK;
}

template <bool CULL, bool COUNT>
static int wf_run(SpkContext *ctx, const WfArgs &a, WfState st, const WfSort &sort, uint32_t sample_count)
{
#ifndef SPK_WF_REFILL
#define SPK_WF_REFILL 8
#endif
    auto extend = wf_extend_kernel<CULL, COUNT, SPK_WF_REFILL>;
    static int blocks_per_sm = 0, sms = 0;
    if (blocks_per_sm == 0)
    {
        SPK_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, extend, kExtendWarps * 32, 0));
        SPK_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
        blocks_per_sm = blocks_per_sm < 1 ? 1 : blocks_per_sm;
        SPK_CUDA_TRY(spk::prefer_small_carveout(extend, sizeof(uint32_t) * kExtendWarps * kLaneSmemWords * 32, blocks_per_sm,
                                                 ctx->opt_carveout));
    }
    const unsigned long long total = (unsigned long long)a.width * a.height * sample_count;
    const uint32_t slots = st.slots;
    const unsigned slot_blocks = (slots + kShadeThreads - 1) / kShadeThreads;
    unsigned extend_blocks = (slots + kExtendWarps * 32 - 1) / (kExtendWarps * 32);
    const unsigned resident = unsigned(blocks_per_sm) * unsigned(sms); // one wave of persistent CTAs
    extend_blocks = extend_blocks < resident ? extend_blocks : resident;
    const unsigned regen_blocks = slot_blocks < unsigned(sms) * 8u ? slot_blocks : unsigned(sms) * 8u;
    cudaStream_t s = ctx->stream;
    wf_init_kernel<<<slot_blocks, kShadeThreads, 0, s>>>(st, total);
    ctx->stats.kernel_launches++;
    // every iteration traces one segment of every live slot; the host looks at the control block once per chunk
    // first look after the fewest iterations the job can take (one segment per path), then every few iterations
    unsigned long long least = (total + slots - 1) / slots + a.max_bounce;
    uint32_t chunk = uint32_t(least < 4096ull ? least : 4096ull);
    const uint32_t chunk_next = ctx->opt_wf_chunk;
    uint32_t cur = 0;
    // ray sorting: three queue buffers rotate. The queue an iteration fills is sorted by ray key into the spare buffer,
    // which wf_extend and the next wf_shade read; the counts in the control block stay indexed by parity.
    uint32_t *q_cur = st.queue[0], *q_nxt = st.queue[1], *q_spare = sort.queue_spare;
    uint32_t queue_bound = slots; // upper bound of a queue's length: all slots until the work runs out, then what is left
    WfCtrl h{};
    for (;; chunk = chunk_next)
    {
        for (uint32_t it = 0; it < chunk; it++)
        {
            st.queue[cur] = q_cur;
            st.queue[cur ^ 1u] = q_nxt;
            // the grid covers what the queue can hold: all slots while new paths start, then (the queue only shrinks) its
            // length at the host's last look
            wf_shade_kernel<<<(queue_bound + kShadeThreads - 1) / kShadeThreads, kShadeThreads, 0, s>>>(ctx->scene, a, st, cur,
                                                                                                         ctx->d_accum);
            wf_regen_kernel<<<regen_blocks, kShadeThreads, 0, s>>>(ctx->camera, a, st, cur ^ 1u);
            wf_advance_kernel<<<1, 1, 0, s>>>(st, cur, ctx->d_counters);
            ctx->stats.kernel_launches += 4;
            if (a.key_bits)
            {
                size_t bytes = sort.tmp_bytes;
                SPK_CUDA_TRY(cub::DeviceRadixSort::SortPairs(sort.tmp, bytes, (const uint32_t *)st.keys, sort.keys_sorted,
                                                             (const uint32_t *)q_nxt, q_spare, queue_bound, 0, sort.end_bit, s));
                uint32_t *t = q_nxt;
                q_nxt = q_spare;
                q_spare = t;
                st.queue[cur ^ 1u] = q_nxt;
            }
            extend<<<extend_blocks, kExtendWarps * 32, 0, s>>>(ctx->scene, st, cur ^ 1u, ctx->d_counters);
            // next iteration: the traced queue is read, the one just consumed is refilled
            uint32_t *t = q_cur;
            q_cur = q_nxt;
            q_nxt = t;
            cur ^= 1u;
        }
        SPK_CUDA_TRY(cudaGetLastError());
        SPK_CUDA_TRY(cudaMemcpyAsync(&h, st.ctrl, sizeof(h), cudaMemcpyDeviceToHost, s));
        SPK_CUDA_TRY(cudaStreamSynchronize(s));
        if (h.work_next >= h.work_total && h.n_queue[cur] == 0)
        {
            break;
        }
        if (h.work_next >= h.work_total && h.n_queue[cur] < queue_bound)
        {
            queue_bound = h.n_queue[cur]; // no new paths: the queue only shrinks from here
        }
    }
    return SPK_OK;
}

int launch_render_wavefront(SpkContext *ctx, const SpkRenderParams *p)
{
    if (p->accum_mode != SPK_ACCUM_SUM)
    {
        set_error("spk_cuda_render: the wavefront tracer adds samples in arbitrary order: SPK_ACCUM_SUM only");
        return SPK_ERR_UNSUPPORTED;
    }
    if ((p->flags >> 8) & 3u)
    {
        set_error("spk_cuda_render: parity probes (debug flags) are a megakernel feature");
        return SPK_ERR_UNSUPPORTED;
    }
    int rc_clear = ensure_accum_cleared(ctx); // samples are added with atomics: a pending reset is cleared now
    if (rc_clear != SPK_OK)
    {
        return rc_clear;
    }
    WfArgs a;
    a.width = ctx->width;
    a.height = ctx->height;
    a.first_sample = p->first_sample;
    a.sample_stride = p->sample_stride ? p->sample_stride : 1u;
    a.max_bounce = p->max_bounce;
    a.output_limit = p->output_limit;
    const unsigned long long total = (unsigned long long)a.width * a.height * p->sample_count;
    uint32_t slots = ctx->opt_wf_slots;
    if (slots > total)
    {
        slots = uint32_t(total);
    }
    slots = (slots + 255u) & ~255u;
    // ray sorting: cells of the scene box (the TLAS root's bounds)
    a.key_bits = ctx->opt_wf_sort;
    for (int k = 0; k < 3; k++)
    {
        const float lo = ctx->tlas_tree.root_bounds[k], hi = ctx->tlas_tree.root_bounds[3 + k];
        a.key_lo[k] = lo;
        a.key_scale[k] = hi > lo ? float(1u << a.key_bits) / (hi - lo) : 0.f;
    }
    if (ctx->scene.num_instances == 0)
    {
        a.key_bits = 0;
    }
    WfState st;
    WfSort sort;
    int rc = wf_reserve(ctx, slots, st, sort, a.key_bits);
    if (rc != SPK_OK)
    {
        return rc;
    }
    const bool exact = p->flags & SPK_FLAG_EXACT_TRAVERSAL, count = p->flags & SPK_FLAG_COUNTERS;
    if (exact && count)
        return wf_run<false, true>(ctx, a, st, sort, p->sample_count);
    if (exact)
        return wf_run<false, false>(ctx, a, st, sort, p->sample_count);
    if (count)
        return wf_run<true, true>(ctx, a, st, sort, p->sample_count);
    return wf_run<true, false>(ctx, a, st, sort, p->sample_count);
}
} // namespace spk
