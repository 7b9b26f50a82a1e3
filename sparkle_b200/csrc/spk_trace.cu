/*
 * spk_trace.cu — closest-hit traversal of a ray batch: SceneRenderProxy::Intersect<false> per ray
 * (libraries/source/renderer/proxy/SceneRenderProxy.cpp:82-126, 299-312). Compiled with -fmad=false.
 *
 * One thread per ray, 128-thread CTAs. Rays are 32 B (two 128-bit loads), hits 20 B. The node-pair and triangle
 * arrays are read through the read-only path with 128-bit loads (4 per inner visit, 3 per triangle).
 * Roofline: HBM/L2-bound pointer chase; algorithmic bytes per ray = 32 + 20 + 64*N_inner + 48*N_tri + 128*N_inst
 * (SURVEY.md 8d), with the N counted by the COUNT=true instantiation of the same code.
 */
#include "spk_context.h"

#include <algorithm>
#include <cstdlib>
#include <mutex>
#include <string>
#include "spk_traverse.cuh"
#include "spk_traverse_warp.cuh"

namespace spk
{
namespace
{
constexpr int kTraceThreads = 128;

template <bool CULL, bool COUNT>
__global__ void __launch_bounds__(kTraceThreads)
    trace_closest_kernel(const DScene sc, const float4 *__restrict__ rays, uint64_t n, float *__restrict__ hits,
                         unsigned long long *__restrict__ counters)
{
    uint64_t i = uint64_t(blockIdx.x) * kTraceThreads + threadIdx.x;
    TraceCounters cnt{0, 0, 0, 0};
    if (i < n)
    {
        float4 ro = __ldg(rays + 2 * i), rd = __ldg(rays + 2 * i + 1);
        uint32_t stack[kThreadStack];
        Hit best;
        trace_closest<CULL, COUNT>(sc, mk(ro.x, ro.y, ro.z), mk(rd.x, rd.y, rd.z), best, stack, cnt);
        float *o = hits + 5 * i;
        o[0] = best.t;
        o[1] = best.u;
        o[2] = best.v;
        uint32_t prim = best.inst == SPK_INVALID_ID ? SPK_INVALID_ID : __ldg(&sc.instances[best.inst].prim);
        o[3] = __uint_as_float(prim);
        o[4] = __uint_as_float(best.face);
    }
    if (COUNT)
    {
        // warp-reduce, one atomic per warp and counter
        uint32_t v[4] = {cnt.node_visits, cnt.tri_tests, cnt.inst_visits, cnt.sphere_tests};
#pragma unroll
        for (int k = 0; k < 4; k++)
        {
            uint32_t s = __reduce_add_sync(0xffffffffu, v[k]);
            if ((threadIdx.x & 31) == 0 && s)
            {
                atomicAdd(counters + 2 + k, (unsigned long long)s);
            }
        }
    }
}

// ---- persistent, warp-converged form (spk_traverse_warp.cuh): the default for ray batches ----
constexpr int kPersistWarps = 4;
#ifndef SPK_TRACE_REFILL
#define SPK_TRACE_REFILL 16
#endif

template <bool CULL, bool COUNT, int VOTE, bool TOP = false>
__global__ void __launch_bounds__(kPersistWarps * 32)
    trace_persistent_kernel(const DScene sc, const float4 *__restrict__ rays, unsigned long long n,
                            float *__restrict__ hits, unsigned long long *__restrict__ counters,
                            unsigned long long *__restrict__ next_ray, uint32_t top_pairs)
{
    __shared__ uint32_t s_lane_words[kPersistWarps * kLaneSmemWords * 32];
    extern __shared__ __align__(16) unsigned char s_top_bytes[]; // TOP: top_pairs x 64 bytes
    TopOfTree top{reinterpret_cast<const float4 *>(s_top_bytes), sc.top_base, TOP ? min(sc.top_count, top_pairs) : 0u};
    if (TOP)
    {
        stage_top_of_tree(sc.pairs, reinterpret_cast<float4 *>(s_top_bytes), top.base, top.count);
    }
    TraceCounters cnt{0, 0, 0, 0};
    auto write_hit = [&](unsigned long long id, const Hit &best) {
        float *o = hits + 5 * id;
        o[0] = best.t;
        o[1] = best.u;
        o[2] = best.v;
        uint32_t prim = best.inst == SPK_INVALID_ID ? SPK_INVALID_ID : __ldg(&sc.instances[best.inst].prim);
        o[3] = __uint_as_float(prim);
        o[4] = __uint_as_float(best.face);
    };
    warp_trace_loop<CULL, COUNT, SPK_TRACE_REFILL, VOTE, TOP>(
        sc, n, next_ray, s_lane_words, cnt, ~0ull,
        [&](unsigned long long id, V3 &o, V3 &d, unsigned long long &cookie) {
            if (sc.num_instances == 0)
            {
                Hit miss{SPK_FLT_MAX, 0.f, 0.f, SPK_INVALID_ID, SPK_INVALID_ID}; // empty scene: every ray misses
                write_hit(id, miss);
                return false;
            }
            float4 ro = __ldg(rays + 2 * id), rd = __ldg(rays + 2 * id + 1);
            o = mk(ro.x, ro.y, ro.z);
            d = mk(rd.x, rd.y, rd.z);
            cookie = id;
            return true;
        },
        write_hit, top);
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
} // namespace

template <bool CULL, bool COUNT, int VOTE, bool TOP = false>
static int launch_persistent(SpkContext *ctx, const float4 *rays, uint64_t n, float *hits)
{
    auto kernel = trace_persistent_kernel<CULL, COUNT, VOTE, TOP>;
    const uint32_t top_pairs = TOP ? uint32_t(ctx->opt_trace_top) : 0u;
    const size_t dynamic = size_t(top_pairs) * 64;
    static int blocks_per_sm = 0, sms = 0;
    static size_t measured_for = ~size_t(0);
    if (measured_for != dynamic)
    {
        if (dynamic > 48 * 1024)
        {
            SPK_CUDA_TRY(cudaFuncSetAttribute(kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)dynamic));
        }
        const size_t per_cta = sizeof(uint32_t) * kPersistWarps * kLaneSmemWords * 32 + dynamic;
        // ask for the carve-out that holds 8 CTAs (the register limit) if the array allows it, then see what fits
        SPK_CUDA_TRY(spk::prefer_small_carveout(kernel, per_cta, 8, ctx->opt_carveout));
        SPK_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&blocks_per_sm, kernel, kPersistWarps * 32, dynamic));
        SPK_CUDA_TRY(cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, ctx->device));
        blocks_per_sm = blocks_per_sm < 1 ? 1 : blocks_per_sm;
        SPK_CUDA_TRY(spk::prefer_small_carveout(kernel, per_cta, blocks_per_sm, ctx->opt_carveout));
        measured_for = dynamic;
    }
    uint64_t want = (n + kPersistWarps * 32 - 1) / (kPersistWarps * 32);
    uint64_t resident = uint64_t(blocks_per_sm) * sms; // one wave of persistent CTAs: a multiple of the SM count
    unsigned grid = unsigned(want < resident ? want : resident);
    unsigned long long *next_ray = ctx->d_counters + 7;
    SPK_CUDA_TRY(cudaMemsetAsync(next_ray, 0, sizeof(unsigned long long), ctx->stream));
    kernel<<<grid, kPersistWarps * 32, dynamic, ctx->stream>>>(ctx->scene, rays, n, hits, ctx->d_counters, next_ray, top_pairs);
    return SPK_OK;
}

int launch_trace_closest(SpkContext *ctx, const void *d_rays, uint64_t n, void *d_hits, uint32_t flags)
{
    if (n == 0)
    {
        return SPK_OK;
    }
    bool exact = flags & SPK_FLAG_EXACT_TRAVERSAL, count = flags & SPK_FLAG_COUNTERS;
    auto *rays = static_cast<const float4 *>(d_rays);
    auto *hits = static_cast<float *>(d_hits);
    int rc = SPK_OK;
    if (flags & SPK_FLAG_THREAD_PER_RAY)
    {
        uint64_t blocks = (n + kTraceThreads - 1) / kTraceThreads;
        if (blocks > 0x7fffffffull)
        {
            set_error("spk_cuda_trace_closest: batch too large for one thread-per-ray launch");
            return SPK_ERR_INVALID_ARGUMENT;
        }
        dim3 grid{unsigned(blocks)}, block{kTraceThreads};
        if (exact && count)
            trace_closest_kernel<false, true><<<grid, block, 0, ctx->stream>>>(ctx->scene, rays, n, hits, ctx->d_counters);
        else if (exact)
            trace_closest_kernel<false, false><<<grid, block, 0, ctx->stream>>>(ctx->scene, rays, n, hits, ctx->d_counters);
        else if (count)
            trace_closest_kernel<true, true><<<grid, block, 0, ctx->stream>>>(ctx->scene, rays, n, hits, ctx->d_counters);
        else
            trace_closest_kernel<true, false><<<grid, block, 0, ctx->stream>>>(ctx->scene, rays, n, hits, ctx->d_counters);
    }
    else if (exact && count)
        rc = launch_persistent<false, true, 8>(ctx, rays, n, hits);
    else if (exact)
        rc = launch_persistent<false, false, 8>(ctx, rays, n, hits);
    else if (count)
        rc = launch_persistent<true, true, 8>(ctx, rays, n, hits);
    else
    {
        if (ctx->opt_trace_top > 0)
        {
            rc = launch_persistent<true, false, 6, true>(ctx, rays, n, hits);
        }
        else
        switch (ctx->opt_vote)
        {
        case 4: rc = launch_persistent<true, false, 4>(ctx, rays, n, hits); break;
        case 8: rc = launch_persistent<true, false, 8>(ctx, rays, n, hits); break;
        case 12: rc = launch_persistent<true, false, 12>(ctx, rays, n, hits); break;
        case 16: rc = launch_persistent<true, false, 16>(ctx, rays, n, hits); break;
        default: rc = launch_persistent<true, false, 6>(ctx, rays, n, hits); break;
        }
    }
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaGetLastError());
    ctx->stats.kernel_launches++;
    ctx->stats.rays += n;
    return SPK_OK;
}
} // namespace spk

using namespace spk;

extern "C"
{
int spk_cuda_trace_closest_device(SpkContext *ctx, const void *d_rays, uint64_t num_rays, void *d_hits, uint32_t flags)
{
    if (!ctx || (num_rays && (!d_rays || !d_hits)))
    {
        set_error("spk_cuda_trace_closest_device: null argument");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (!ctx->has_scene)
    {
        set_error("spk_cuda_trace_closest: no scene uploaded");
        return SPK_ERR_NO_SCENE;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev0, ctx->stream));
    int rc = launch_trace_closest(ctx, d_rays, num_rays, d_hits, flags);
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev1, ctx->stream));
    return SPK_OK;
}

int spk_cuda_trace_closest(SpkContext *ctx, const SpkRay *rays, uint64_t num_rays, SpkHit *hits, uint32_t flags)
{
    if (!ctx || (num_rays && (!rays || !hits)))
    {
        set_error("spk_cuda_trace_closest: null argument");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (!ctx->has_scene)
    {
        set_error("spk_cuda_trace_closest: no scene uploaded");
        return SPK_ERR_NO_SCENE;
    }
    if (num_rays == 0)
    {
        return SPK_OK;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    if (num_rays > ctx->ray_capacity)
    {
        cudaFree(ctx->d_rays);
        cudaFree(ctx->d_hits);
        ctx->d_rays = ctx->d_hits = nullptr;
        ctx->ray_capacity = 0;
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_rays, num_rays * sizeof(SpkRay)));
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_hits, num_rays * sizeof(SpkHit)));
        ctx->ray_capacity = num_rays;
    }
    // Page-locked host buffers (spk_cuda_host_alloc, cudaHostAlloc, pinned torch tensors): a three-stage pipeline over
    // pieces of 2 Mi rays - copy-in, traversal and copy-out on three streams - so the two PCIe directions and the kernel
    // overlap; the call is then bound by the slower direction (32 B in per ray) instead of the sum of all three.
    auto is_pinned = [](const void *p) {
        cudaPointerAttributes attr{};
        const bool yes = cudaPointerGetAttributes(&attr, p) == cudaSuccess && attr.type == cudaMemoryTypeHost;
        cudaGetLastError();
        return yes;
    };
    constexpr uint64_t kPiece = 2u << 20;
    if (num_rays >= 2 * kPiece && is_pinned(rays) && is_pinned(hits))
    {
        if (!ctx->copy_in_stream)
        {
            SPK_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->copy_in_stream, cudaStreamNonBlocking));
            SPK_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->copy_out_stream, cudaStreamNonBlocking));
            for (int k = 0; k < 8; k++)
            {
                SPK_CUDA_TRY(cudaEventCreateWithFlags(&ctx->pipe_ev[k], cudaEventDisableTiming));
                SPK_CUDA_TRY(cudaEventCreateWithFlags(&ctx->pipe_ev[8 + k], cudaEventDisableTiming));
            }
        }
        // the copy-in stream starts where the context's stream is now (earlier work may still be using the buffers)
        SPK_CUDA_TRY(cudaEventRecord(ctx->pipe_ev[0], ctx->stream));
        SPK_CUDA_TRY(cudaStreamWaitEvent(ctx->copy_in_stream, ctx->pipe_ev[0], 0));
        SPK_CUDA_TRY(cudaEventRecord(ctx->ev0, ctx->stream));
        uint64_t piece = 0;
        for (uint64_t at = 0; at < num_rays; at += kPiece, piece++)
        {
            const uint64_t n = std::min<uint64_t>(kPiece, num_rays - at);
            cudaEvent_t in_done = ctx->pipe_ev[piece % 8], traced = ctx->pipe_ev[8 + piece % 8];
            SPK_CUDA_TRY(cudaMemcpyAsync(static_cast<SpkRay *>(ctx->d_rays) + at, rays + at, n * sizeof(SpkRay),
                                         cudaMemcpyHostToDevice, ctx->copy_in_stream));
            SPK_CUDA_TRY(cudaEventRecord(in_done, ctx->copy_in_stream));
            SPK_CUDA_TRY(cudaStreamWaitEvent(ctx->stream, in_done, 0));
            int rc = launch_trace_closest(ctx, static_cast<SpkRay *>(ctx->d_rays) + at, n, static_cast<SpkHit *>(ctx->d_hits) + at,
                                          flags);
            if (rc != SPK_OK)
            {
                return rc;
            }
            SPK_CUDA_TRY(cudaEventRecord(traced, ctx->stream));
            SPK_CUDA_TRY(cudaStreamWaitEvent(ctx->copy_out_stream, traced, 0));
            SPK_CUDA_TRY(cudaMemcpyAsync(hits + at, static_cast<SpkHit *>(ctx->d_hits) + at, n * sizeof(SpkHit),
                                         cudaMemcpyDeviceToHost, ctx->copy_out_stream));
        }
        SPK_CUDA_TRY(cudaEventRecord(ctx->ev1, ctx->stream));
        SPK_CUDA_TRY(cudaStreamSynchronize(ctx->copy_out_stream));
        SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        return SPK_OK;
    }
    SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_rays, rays, num_rays * sizeof(SpkRay), cudaMemcpyHostToDevice, ctx->stream));
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev0, ctx->stream));
    int rc = launch_trace_closest(ctx, ctx->d_rays, num_rays, ctx->d_hits, flags);
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev1, ctx->stream));
    SPK_CUDA_TRY(cudaMemcpyAsync(hits, ctx->d_hits, num_rays * sizeof(SpkHit), cudaMemcpyDeviceToHost, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return SPK_OK;
}

} // extern "C"

// ---- spk_cuda_trace_closest_noalloc: the same ray batch call with NO allocation inside ------------------------------
// spk_cuda_trace_closest stages the host batch in ctx->d_rays / ctx->d_hits (cudaMalloc'd on demand) and hands out rays
// through a ticket counter. Here all three live in ONE file-scope device array that every GPU loading the library holds
// until exit: [0, 256) the ticket (must be zero at launch: cleared on the stream before every launch), then the staged
// rays (32 B each), then the hits (20 B each). The array keeps the previous call's rays and hits; they are overwritten
// before they are read (rays by the copy-in, hits by the kernel: every ray writes its hit, misses included).
namespace spk
{
namespace
{
constexpr uint64_t kNoAllocMaxRays = SPK_NOALLOC_MAX_RAYS;
constexpr size_t kNoAllocTicketBytes = 256;
constexpr size_t kNoAllocRaysAt = kNoAllocTicketBytes;
constexpr size_t kNoAllocHitsAt = kNoAllocRaysAt + kNoAllocMaxRays * sizeof(SpkRay);
constexpr size_t kNoAllocBytes = kNoAllocHitsAt + kNoAllocMaxRays * sizeof(SpkHit);
static_assert(sizeof(SpkRay) == 32 && sizeof(SpkHit) == 20, "scratch layout follows the C ABI's ray and hit records");
static_assert(kNoAllocHitsAt % 256 == 0, "hits start on a 256-byte boundary");

static __device__ __align__(256) unsigned char g_noalloc_scratch[kNoAllocBytes];

constexpr int kNoAllocMaxDevices = 64;
unsigned char *g_noalloc_ptr[kNoAllocMaxDevices];    // device address of the array on each device, looked up once
int g_noalloc_blocks_per_sm[2], g_noalloc_sms[kNoAllocMaxDevices];
std::mutex g_noalloc_mutex;

// the array's name in host code is a host-side stand-in: the kernel needs what cudaGetSymbolAddress returns on the device
// that is current
int noalloc_scratch(SpkContext *ctx, unsigned char **out)
{
    if (ctx->device < 0 || ctx->device >= kNoAllocMaxDevices)
    {
        set_error("spk_cuda_trace_closest_noalloc: device ordinal beyond 63");
        return SPK_ERR_UNSUPPORTED;
    }
    std::lock_guard<std::mutex> lock(g_noalloc_mutex);
    if (!g_noalloc_ptr[ctx->device])
    {
        void *p = nullptr;
        SPK_CUDA_TRY(cudaGetSymbolAddress(&p, g_noalloc_scratch));
        SPK_CUDA_TRY(cudaDeviceGetAttribute(&g_noalloc_sms[ctx->device], cudaDevAttrMultiProcessorCount, ctx->device));
        g_noalloc_ptr[ctx->device] = static_cast<unsigned char *>(p);
    }
    *out = g_noalloc_ptr[ctx->device];
    return SPK_OK;
}

template <bool CULL>
int launch_noalloc(SpkContext *ctx, unsigned char *scratch, uint64_t n)
{
    auto kernel = trace_persistent_kernel<CULL, false, 6, false>;
    int &blocks_per_sm = g_noalloc_blocks_per_sm[CULL ? 1 : 0];
    if (blocks_per_sm == 0)
    {
        int b = 0;
        SPK_CUDA_TRY(cudaOccupancyMaxActiveBlocksPerMultiprocessor(&b, kernel, kPersistWarps * 32, 0));
        blocks_per_sm = b < 1 ? 1 : b;
    }
    const uint64_t want = (n + kPersistWarps * 32 - 1) / (kPersistWarps * 32);
    const uint64_t resident = uint64_t(blocks_per_sm) * g_noalloc_sms[ctx->device];
    const unsigned grid = unsigned(want < resident ? want : resident);
    auto *ticket = reinterpret_cast<unsigned long long *>(scratch);
    SPK_CUDA_TRY(cudaMemsetAsync(ticket, 0, kNoAllocTicketBytes, ctx->stream));
    kernel<<<grid, kPersistWarps * 32, 0, ctx->stream>>>(ctx->scene, reinterpret_cast<const float4 *>(scratch + kNoAllocRaysAt), n,
                                                        reinterpret_cast<float *>(scratch + kNoAllocHitsAt), ctx->d_counters,
                                                        ticket, 0u);
    SPK_CUDA_TRY(cudaGetLastError());
    return SPK_OK;
}
} // namespace
} // namespace spk

extern "C"
{
int spk_cuda_trace_closest_noalloc(SpkContext *ctx, const SpkRay *rays, uint64_t num_rays, SpkHit *hits, uint32_t flags)
{
    if (!ctx || (num_rays && (!rays || !hits)))
    {
        set_error("spk_cuda_trace_closest_noalloc: null argument");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (flags & ~uint32_t(SPK_FLAG_EXACT_TRAVERSAL))
    {
        set_error("spk_cuda_trace_closest_noalloc: only SPK_FLAG_EXACT_TRAVERSAL is accepted");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (num_rays > kNoAllocMaxRays)
    {
        set_error("spk_cuda_trace_closest_noalloc: " + std::to_string(num_rays) + " rays exceed the static scratch's limit of " +
                  std::to_string(kNoAllocMaxRays) + " (SPK_NOALLOC_MAX_RAYS); split the batch or use spk_cuda_trace_closest");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    if (!ctx->has_scene)
    {
        set_error("spk_cuda_trace_closest_noalloc: no scene uploaded");
        return SPK_ERR_NO_SCENE;
    }
    if (num_rays == 0)
    {
        return SPK_OK;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    unsigned char *scratch = nullptr;
    int rc = noalloc_scratch(ctx, &scratch);
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaMemcpyAsync(scratch + kNoAllocRaysAt, rays, num_rays * sizeof(SpkRay), cudaMemcpyHostToDevice, ctx->stream));
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev0, ctx->stream));
    rc = (flags & SPK_FLAG_EXACT_TRAVERSAL) ? launch_noalloc<false>(ctx, scratch, num_rays)
                                            : launch_noalloc<true>(ctx, scratch, num_rays);
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev1, ctx->stream));
    SPK_CUDA_TRY(cudaMemcpyAsync(hits, scratch + kNoAllocHitsAt, num_rays * sizeof(SpkHit), cudaMemcpyDeviceToHost, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->stats.kernel_launches++;
    ctx->stats.rays += num_rays;
    return SPK_OK;
}

int spk_cuda_get_stats(SpkContext *ctx, SpkStats *out)
{
    if (!ctx || !out)
    {
        set_error("spk_cuda_get_stats: null argument");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    unsigned long long c[8];
    SPK_CUDA_TRY(cudaMemcpy(c, ctx->d_counters, sizeof(c), cudaMemcpyDeviceToHost));
    ctx->stats.paths = c[0];
    ctx->stats.node_visits = c[2];
    ctx->stats.tri_tests = c[3];
    ctx->stats.inst_visits = c[4];
    ctx->stats.sphere_tests = c[5];
    float ms = 0.f;
    if (cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1) == cudaSuccess)
    {
        ctx->stats.last_ms = ms;
    }
    else
    {
        (void)cudaGetLastError();
    }
    *out = ctx->stats;
    out->rays += c[1]; // path segments the path tracer counted on the device
    return SPK_OK;
}

int spk_cuda_reset_stats(SpkContext *ctx)
{
    if (!ctx)
    {
        set_error("null context");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
    uint64_t launches = ctx->stats.kernel_launches;
    ctx->stats = SpkStats{};
    ctx->stats.kernel_launches = launches;
    return SPK_OK;
}

int spk_cuda_synchronize(SpkContext *ctx)
{
    if (!ctx)
    {
        set_error("null context");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    return SPK_OK;
}
} // extern "C"
