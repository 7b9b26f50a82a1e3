/*
 * spk_group.cu — multi-GPU behind the C ABI (north_star (d)): the accumulators of G contexts, each holding
 * an interleaved share of the samples, are combined and tone-mapped by ONE kernel per GPU that does the collective and
 * the pass that follows it together:
 *
 *     reduce-scatter (rows)  +  divide / ACES / pack (ToneMappingPass, CPURenderer.cpp:554-576)  +  gather to the root
 *
 * Rank r owns image rows [r*H/G, (r+1)*H/G). Its kernel reads those rows from ALL G accumulators over NVLink peer memory
 * (plain 16-byte loads from the peers' HBM through NVSwitch), sums them in rank order, tone-maps, and stores the packed
 * pixels straight into the ROOT's output buffer (8 bytes per pixel over NVLink instead of 16 in an ncclReduce followed by
 * a root-only resolve); the root then makes one DMA to the caller's host buffer. Every link carries 1/G of the frame in
 * each direction, all GPUs tone-map in parallel, nothing is staged. Round 1 did this as ncclReduce(root 0) from Python +
 * a rank-0 resolve: the fixed per-frame part that kept end-to-end scaling at 6.8x on 8 GPUs.
 *
 * Two ways to form a group, one kernel:
 *   spk_cuda_group_create_local  all contexts in this process (CudaRenderer with `gpus` > 1): peers are plain device
 *                                pointers (cudaDeviceEnablePeerAccess), ordering between the GPUs is stream events.
 *   spk_cuda_group_create        one process per GPU (torchrun): peers are CUDA IPC mappings of the other ranks'
 *                                accumulators; ordering is a flag protocol in device memory (arrive: "my accumulator is
 *                                complete", depart: "I have read yours and my stores to the root have landed") with
 *                                release/acquire at system scope. Every rank must make the same resolve calls (it is a
 *                                collective); one rank per GPU, as the flags are spun on.
 * The sum is ((a_0 + a_1) + a_2) + ... in rank order whatever the number of GPUs: deterministic, and equal bit for bit
 * to adding the per-rank accumulators in that order on the host (tests/test_multi_gpu_gpu.py).
 */
#include "spk_context.h"
#include "spk_shade.cuh"

#include <cstdio>
#include <cstring>
#include <cuda_fp16.h>
#include <unistd.h>
#include <vector>

namespace spk
{
namespace
{
constexpr uint32_t kMaxRanks = 16;
constexpr uint32_t kMagic = 0x53504B47u; // "SPKG"

// per-context block peers can see: flags first, the root's gathered output after them
struct SyncBlock
{
    uint32_t arrive[kMaxRanks]; // arrive[r] = e: rank r's accumulator holds every sample of frame e
    uint32_t depart[kMaxRanks]; // depart[r] = e: rank r has read this rank's accumulator and stored its rows at the root
    uint32_t blocks_done;       // blocks of the local kernel that finished their rows (reset by the last one)
    uint32_t timed_out;         // set when a bounded spin gave up (a rank did not join): reported by the host
    uint32_t pad[30];
};
static_assert(sizeof(SyncBlock) == 256, "sync block");

struct PeerHandleData
{
    uint32_t magic, abi;
    int32_t device;
    int32_t pid;
    uint32_t width, height;
    uint64_t host_id;
    cudaIpcMemHandle_t accum;
    cudaIpcMemHandle_t block;
};
static_assert(sizeof(PeerHandleData) <= SPK_PEER_HANDLE_BYTES, "SpkPeerHandle holds the exported handles");

struct GroupArgs
{
    const float4 *accum[kMaxRanks]; // every rank's accumulator (own: local pointer)
    SyncBlock *sync[kMaxRanks];     // every rank's sync block (flag protocol only)
    void *out_root;                 // the root's output buffer
    uint32_t rank, world;
    uint32_t width, height, row_begin, row_end;
    uint32_t format;
    float divisor, exposure;
    uint32_t epoch; // 0: ordering is done with stream events (local group), no flags
};

__device__ __forceinline__ void st_release_sys(uint32_t *p, uint32_t v)
{
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(p), "r"(v) : "memory");
}
__device__ __forceinline__ uint32_t ld_acquire_sys(const uint32_t *p)
{
    uint32_t v;
    asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(p) : "memory");
    return v;
}
// peers' accumulators are read once and change every frame: no caching at any level on this GPU
__device__ __forceinline__ float4 ld_peer(const float4 *p)
{
    float4 v;
    asm volatile("ld.volatile.global.v4.f32 {%0,%1,%2,%3}, [%4];" : "=f"(v.x), "=f"(v.y), "=f"(v.z), "=f"(v.w) : "l"(p));
    return v;
}

// a rank that never joins must not hang the GPU: ~8 s at 2 GHz, then the frame is reported as failed
constexpr long long kSpinLimit = 16000000000ll;

__device__ bool wait_flags(const uint32_t *flags, uint32_t world, uint32_t epoch, SyncBlock *mine)
{
    // threads 0..world-1 each watch one peer's flag
    bool ok = true;
    if (threadIdx.x < world)
    {
        const long long t0 = clock64();
        while ((int32_t)(ld_acquire_sys(flags + threadIdx.x) - epoch) < 0)
        {
            if (clock64() - t0 > kSpinLimit)
            {
                mine->timed_out = 1u;
                ok = false;
                break;
            }
            __nanosleep(200);
        }
    }
    return __syncthreads_and(ok);
}

__device__ __forceinline__ float aces_film(float c)
{
    // ACESFilm (CPURenderer.cpp:554-563), the same expression as resolve_kernel
    float r = (c * (c * 2.51f + 0.03f)) / (c * (c * 2.43f + 0.59f) + 0.14f);
    return fminf(fmaxf(r, 0.f), 1.f);
}

__global__ void __launch_bounds__(256) group_reduce_resolve_kernel(const GroupArgs g)
{
    SyncBlock *mine = g.sync[g.rank];
    if (g.epoch)
    {
        // arrive: this kernel runs after the render on the same stream, so the local accumulator is complete
        if (blockIdx.x == 0 && threadIdx.x < g.world)
        {
            st_release_sys(&g.sync[threadIdx.x]->arrive[g.rank], g.epoch);
        }
        // every block waits for all accumulators before reading them (the flags are local memory: cheap to poll)
        if (!wait_flags(mine->arrive, g.world, g.epoch, mine))
        {
            return;
        }
    }
    const uint32_t rows = g.row_end - g.row_begin;
    const uint32_t idx = blockIdx.x * blockDim.x + threadIdx.x;
    if (idx < rows * g.width)
    {
        const uint32_t j = g.row_begin + idx / g.width, i = idx % g.width;
        const size_t px = size_t(j) * g.width + i;
        float4 p = g.rank == 0 ? g.accum[0][px] : ld_peer(g.accum[0] + px);
        for (uint32_t r = 1; r < g.world; r++)
        {
            const float4 q = r == g.rank ? g.accum[r][px] : ld_peer(g.accum[r] + px);
            p = make_float4(p.x + q.x, p.y + q.y, p.z + q.z, p.w + q.w);
        }
        if (g.divisor != 1.f)
        {
            p = make_float4(p.x / g.divisor, p.y / g.divisor, p.z / g.divisor, p.w / g.divisor);
        }
        if (g.format == SPK_RESOLVE_LINEAR_F32)
        {
            static_cast<float4 *>(g.out_root)[px] = p;
        }
        else
        {
            const float r = aces_film(p.x * g.exposure), gr = aces_film(p.y * g.exposure), b = aces_film(p.z * g.exposure);
            const size_t o = size_t(g.height - 1 - j) * g.width + i; // image.SetPixel(i, H-1-j, ...)
            if (g.format == SPK_RESOLVE_TONEMAP_F16)
            {
                __half2 lo = __floats2half2_rn(r, gr), hi = __floats2half2_rn(b, 1.f);
                uint2 raw;
                raw.x = *reinterpret_cast<uint32_t *>(&lo);
                raw.y = *reinterpret_cast<uint32_t *>(&hi);
                static_cast<uint2 *>(g.out_root)[o] = raw;
            }
            else
            {
                uchar4 c;
                c.x = (unsigned char)(ref_pow(r, 0.45f) * 255.f);
                c.y = (unsigned char)(ref_pow(gr, 0.45f) * 255.f);
                c.z = (unsigned char)(ref_pow(b, 0.45f) * 255.f);
                c.w = 255;
                static_cast<uchar4 *>(g.out_root)[o] = c;
            }
        }
    }
    if (!g.epoch)
    {
        return;
    }
    // depart: the LAST block of this rank tells every peer that this rank is done with their accumulators and that
    // its rows are at the root, then waits for the same from all of them: when the kernel ends, the root's buffer is
    // complete and this rank's accumulator may be overwritten by the next frame.
    __shared__ uint32_t s_last;
    __threadfence_system();
    __syncthreads();
    if (threadIdx.x == 0)
    {
        s_last = atomicAdd(&mine->blocks_done, 1u) == gridDim.x - 1 ? 1u : 0u;
    }
    __syncthreads();
    if (!s_last)
    {
        return;
    }
    if (threadIdx.x == 0)
    {
        mine->blocks_done = 0u;
    }
    __threadfence_system();
    if (threadIdx.x < g.world)
    {
        st_release_sys(&g.sync[threadIdx.x]->depart[g.rank], g.epoch);
    }
    wait_flags(mine->depart, g.world, g.epoch, mine);
}
} // namespace
} // namespace spk

using namespace spk;

struct SpkGroup
{
    bool local = false;
    uint32_t world = 0, rank = 0; // rank: IPC groups only
    std::vector<SpkContext *> ctxs; // local: all members; IPC: the one member of this process
    // IPC: mapped peer pointers (own entries are the local allocations)
    std::vector<void *> mapped;     // what cudaIpcCloseMemHandle must be called on
    const float4 *accum[kMaxRanks] = {};
    SyncBlock *sync[kMaxRanks] = {};
    void *out_root = nullptr;       // IPC: the root's output buffer as seen from this rank
    uint32_t epoch = 0;
    bool async_nonroot = false; // IPC, ranks > 0: do not block in resolve (spk_cuda_group_set_async)
    // local: ordering events, one pair per member
    std::vector<cudaEvent_t> ev_ready, ev_done;
};

namespace
{
int invalid_group(const char *msg)
{
    set_error(msg);
    return SPK_ERR_INVALID_ARGUMENT;
}

size_t out_bytes_max(const SpkContext *ctx)
{
    return size_t(ctx->width) * ctx->height * 16;
}

// the block peers can map: SyncBlock, then the gathered output (only the root's is written)
int ensure_group_block(SpkContext *ctx)
{
    if (!ctx->d_group_block)
    {
        SPK_CUDA_TRY(cudaSetDevice(ctx->device));
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_group_block, sizeof(SyncBlock) + out_bytes_max(ctx)));
        SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_group_block, 0, sizeof(SyncBlock), ctx->stream));
        SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    }
    return SPK_OK;
}

uint64_t host_identity()
{
    // boot id distinguishes hosts: CUDA IPC handles only open on the machine that made them
    uint64_t h = 1469598103934665603ull;
    FILE *f = fopen("/proc/sys/kernel/random/boot_id", "r");
    if (f)
    {
        int c;
        while ((c = fgetc(f)) != EOF)
        {
            h = (h ^ uint64_t(c)) * 1099511628211ull;
        }
        fclose(f);
    }
    return h;
}

void fill_args(GroupArgs &a, const SpkGroup *g, const SpkContext *ctx, uint32_t rank, uint32_t format, uint32_t total_samples,
               uint32_t epoch)
{
    memset(&a, 0, sizeof(a));
    for (uint32_t r = 0; r < g->world; r++)
    {
        a.accum[r] = g->accum[r];
        a.sync[r] = g->sync[r];
    }
    a.out_root = g->out_root;
    a.rank = rank;
    a.world = g->world;
    a.width = ctx->width;
    a.height = ctx->height;
    a.row_begin = uint32_t(uint64_t(ctx->height) * rank / g->world);
    a.row_end = uint32_t(uint64_t(ctx->height) * (rank + 1) / g->world);
    a.format = format;
    a.divisor = ctx->accum_mode == SPK_ACCUM_SUM ? (float)(total_samples ? total_samples : 1u) : 1.f;
    a.exposure = ctx->has_camera ? ctx->camera.exposure : 1.f;
    a.epoch = epoch;
}

int launch_group_kernel(SpkContext *ctx, const GroupArgs &a)
{
    const uint32_t pixels = (a.row_end - a.row_begin) * a.width;
    // at least one block even for an empty slice: the flag protocol needs every rank's arrive and depart
    const unsigned blocks = pixels ? (pixels + 255u) / 256u : 1u;
    group_reduce_resolve_kernel<<<blocks, 256, 0, ctx->stream>>>(a);
    SPK_CUDA_TRY(cudaGetLastError());
    ctx->stats.kernel_launches++;
    return SPK_OK;
}
} // namespace

extern "C"
{
int spk_cuda_group_export(SpkContext *ctx, SpkPeerHandle *out)
{
    if (!ctx || !out)
    {
        return invalid_group("spk_cuda_group_export: null argument");
    }
    int rc = ensure_group_block(ctx);
    if (rc != SPK_OK)
    {
        return rc;
    }
    PeerHandleData h{};
    h.magic = kMagic;
    h.abi = SPK_ABI_VERSION;
    h.device = ctx->device;
    h.pid = (int32_t)getpid();
    h.width = ctx->width;
    h.height = ctx->height;
    h.host_id = host_identity();
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    // a new group counts its frames from 1 again: the flags a previous group of this context left behind are cleared HERE,
    // before the handle can reach any peer (the exchange of the handles is the point every rank passes before it can
    // create its group and signal anybody)
    SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_group_block, 0, sizeof(SyncBlock), ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    SPK_CUDA_TRY(cudaIpcGetMemHandle(&h.accum, ctx->d_accum));
    SPK_CUDA_TRY(cudaIpcGetMemHandle(&h.block, ctx->d_group_block));
    memset(out, 0, sizeof(*out));
    memcpy(out->bytes, &h, sizeof(h));
    return SPK_OK;
}

int spk_cuda_group_create(SpkContext *ctx, uint32_t rank, uint32_t world, const SpkPeerHandle *handles, SpkGroup **out)
{
    if (!ctx || !handles || !out || world == 0 || world > kMaxRanks || rank >= world)
    {
        return invalid_group("spk_cuda_group_create: null argument or bad rank / world (at most 16 ranks)");
    }
    *out = nullptr;
    int rc = ensure_group_block(ctx);
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    auto *g = new (std::nothrow) SpkGroup;
    if (!g)
    {
        return SPK_ERR_OUT_OF_MEMORY;
    }
    g->world = world;
    g->rank = rank;
    g->ctxs.push_back(ctx);
    const uint64_t here = host_identity();
    for (uint32_t r = 0; r < world; r++)
    {
        PeerHandleData h;
        memcpy(&h, handles[r].bytes, sizeof(h));
        if (h.magic != kMagic || h.abi != SPK_ABI_VERSION || h.width != ctx->width || h.height != ctx->height || h.host_id != here)
        {
            spk_cuda_group_destroy(g);
            return invalid_group("spk_cuda_group_create: a peer handle is not from spk_cuda_group_export of a context of the "
                                 "same resolution on this host");
        }
        if (r == rank)
        {
            g->accum[r] = ctx->d_accum;
            g->sync[r] = static_cast<SyncBlock *>(ctx->d_group_block);
            continue;
        }
        if (h.pid == (int32_t)getpid())
        {
            spk_cuda_group_destroy(g);
            return invalid_group("spk_cuda_group_create: peers of the same process form a group with "
                                 "spk_cuda_group_create_local");
        }
        void *pa = nullptr, *pb = nullptr;
        cudaError_t e = cudaIpcOpenMemHandle(&pa, h.accum, cudaIpcMemLazyEnablePeerAccess);
        if (e == cudaSuccess)
        {
            g->mapped.push_back(pa);
            e = cudaIpcOpenMemHandle(&pb, h.block, cudaIpcMemLazyEnablePeerAccess);
        }
        if (e != cudaSuccess)
        {
            spk_cuda_group_destroy(g);
            return cuda_fail(e, "spk_cuda_group_create: cudaIpcOpenMemHandle (peer access between the GPUs is required)");
        }
        g->mapped.push_back(pb);
        g->accum[r] = static_cast<const float4 *>(pa);
        g->sync[r] = static_cast<SyncBlock *>(pb);
    }
    g->out_root = reinterpret_cast<uint8_t *>(g->sync[0]) + sizeof(SyncBlock);
    *out = g;
    return SPK_OK;
}

int spk_cuda_group_create_local(SpkContext *const *ctxs, uint32_t n, SpkGroup **out)
{
    if (!ctxs || !out || n == 0 || n > kMaxRanks)
    {
        return invalid_group("spk_cuda_group_create_local: null argument or more than 16 contexts");
    }
    *out = nullptr;
    for (uint32_t r = 0; r < n; r++)
    {
        if (!ctxs[r] || ctxs[r]->width != ctxs[0]->width || ctxs[r]->height != ctxs[0]->height)
        {
            return invalid_group("spk_cuda_group_create_local: contexts must share one resolution");
        }
    }
    auto *g = new (std::nothrow) SpkGroup;
    if (!g)
    {
        return SPK_ERR_OUT_OF_MEMORY;
    }
    g->local = true;
    g->world = n;
    g->ctxs.assign(ctxs, ctxs + n);
    auto fail = [&](int rc) {
        const std::string why = spk_cuda_last_error();
        spk_cuda_group_destroy(g);
        set_error(why);
        return rc;
    };
    int rc = ensure_group_block(ctxs[0]);
    if (rc != SPK_OK)
    {
        return fail(rc);
    }
    for (uint32_t r = 0; r < n; r++)
    {
        g->accum[r] = ctxs[r]->d_accum;
        g->sync[r] = nullptr;
        cudaError_t e = cudaSetDevice(ctxs[r]->device);
        for (uint32_t q = 0; q < n && e == cudaSuccess; q++)
        {
            if (ctxs[q]->device == ctxs[r]->device)
            {
                continue;
            }
            int can = 0;
            e = cudaDeviceCanAccessPeer(&can, ctxs[r]->device, ctxs[q]->device);
            if (e == cudaSuccess && !can)
            {
                set_error("spk_cuda_group_create_local: no peer access between the devices of the group");
                return fail(SPK_ERR_UNSUPPORTED);
            }
            if (e == cudaSuccess)
            {
                e = cudaDeviceEnablePeerAccess(ctxs[q]->device, 0);
                if (e == cudaErrorPeerAccessAlreadyEnabled)
                {
                    cudaGetLastError();
                    e = cudaSuccess;
                }
            }
        }
        cudaEvent_t a = nullptr, b = nullptr;
        if (e == cudaSuccess)
        {
            e = cudaEventCreateWithFlags(&a, cudaEventDisableTiming);
        }
        if (e == cudaSuccess)
        {
            e = cudaEventCreateWithFlags(&b, cudaEventDisableTiming);
        }
        g->ev_ready.push_back(a);
        g->ev_done.push_back(b);
        if (e != cudaSuccess)
        {
            return fail(cuda_fail(e, "spk_cuda_group_create_local"));
        }
    }
    g->out_root = static_cast<uint8_t *>(ctxs[0]->d_group_block) + sizeof(SyncBlock);
    *out = g;
    return SPK_OK;
}

int spk_cuda_group_resolve(SpkGroup *g, uint32_t format, uint32_t total_samples, void *host_out, size_t host_bytes)
{
    if (!g || format > SPK_RESOLVE_TONEMAP_SRGB8)
    {
        return invalid_group("spk_cuda_group_resolve: null group or unknown format");
    }
    SpkContext *root = g->ctxs[0];
    const size_t pixels = size_t(root->width) * root->height;
    const size_t need = pixels * (format == SPK_RESOLVE_LINEAR_F32 ? 16 : format == SPK_RESOLVE_TONEMAP_F16 ? 8 : 4);
    const bool is_root = g->local || g->rank == 0;
    // host_out == NULL: the frame stays in the root's device buffer and nothing blocks (device-resident pipelines, timing)
    if (is_root && host_out && host_bytes < need)
    {
        return invalid_group("spk_cuda_group_resolve: the host buffer is smaller than the frame");
    }
    int rc;
    if (g->local)
    {
        // every accumulator complete -> every kernel; every kernel done -> the root's copy and everybody's next frame
        for (uint32_t r = 0; r < g->world; r++)
        {
            SpkContext *c = g->ctxs[r];
            SPK_CUDA_TRY(cudaSetDevice(c->device));
            if ((rc = ensure_accum_cleared(c)) != SPK_OK)
            {
                return rc;
            }
            SPK_CUDA_TRY(cudaEventRecord(g->ev_ready[r], c->stream));
        }
        for (uint32_t r = 0; r < g->world; r++)
        {
            SpkContext *c = g->ctxs[r];
            SPK_CUDA_TRY(cudaSetDevice(c->device));
            for (uint32_t q = 0; q < g->world; q++)
            {
                if (q != r)
                {
                    SPK_CUDA_TRY(cudaStreamWaitEvent(c->stream, g->ev_ready[q], 0));
                }
            }
            GroupArgs a;
            fill_args(a, g, c, r, format, total_samples, 0u);
            a.divisor = root->accum_mode == SPK_ACCUM_SUM ? (float)(total_samples ? total_samples : 1u) : 1.f;
            a.exposure = root->has_camera ? root->camera.exposure : 1.f;
            if ((rc = launch_group_kernel(c, a)) != SPK_OK)
            {
                return rc;
            }
            SPK_CUDA_TRY(cudaEventRecord(g->ev_done[r], c->stream));
        }
        for (uint32_t r = 0; r < g->world; r++)
        {
            SpkContext *c = g->ctxs[r];
            SPK_CUDA_TRY(cudaSetDevice(c->device));
            for (uint32_t q = 0; q < g->world; q++)
            {
                if (q != r)
                {
                    SPK_CUDA_TRY(cudaStreamWaitEvent(c->stream, g->ev_done[q], 0));
                }
            }
        }
        SPK_CUDA_TRY(cudaSetDevice(root->device));
        return host_out ? copy_to_host(root, host_out, g->out_root, need) : SPK_OK;
    }
    SpkContext *c = g->ctxs[0];
    SPK_CUDA_TRY(cudaSetDevice(c->device));
    if ((rc = ensure_accum_cleared(c)) != SPK_OK)
    {
        return rc;
    }
    g->epoch++;
    if (g->epoch == 0)
    {
        g->epoch = 1;
    }
    GroupArgs a;
    fill_args(a, g, c, g->rank, format, total_samples, g->epoch);
    if ((rc = launch_group_kernel(c, a)) != SPK_OK)
    {
        return rc;
    }
    if (is_root && !host_out)
    {
        return SPK_OK; // asynchronous: a rank that never joined is reported by the next blocking call
    }
    if (is_root)
    {
        if ((rc = copy_to_host(c, host_out, g->out_root, need)) != SPK_OK)
        {
            return rc;
        }
    }
    else if (host_bytes == 0 && host_out == nullptr && g->async_nonroot)
    {
        return SPK_OK;
    }
    else
    {
        SPK_CUDA_TRY(cudaStreamSynchronize(c->stream));
    }
    uint32_t timed_out = 0;
    SPK_CUDA_TRY(cudaMemcpyAsync(&timed_out, &static_cast<SyncBlock *>(c->d_group_block)->timed_out, 4, cudaMemcpyDeviceToHost,
                                 c->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(c->stream));
    if (timed_out)
    {
        set_error("spk_cuda_group_resolve: a rank of the group did not join the resolve (timed out waiting for its flag)");
        return SPK_ERR_CUDA;
    }
    return SPK_OK;
}

int spk_cuda_group_set_async(SpkGroup *g, uint32_t enabled)
{
    if (!g)
    {
        return invalid_group("spk_cuda_group_set_async: null group");
    }
    g->async_nonroot = enabled != 0;
    return SPK_OK;
}

int spk_cuda_group_destroy(SpkGroup *g)
{
    if (!g)
    {
        return SPK_OK;
    }
    if (!g->ctxs.empty() && g->ctxs[0])
    {
        cudaSetDevice(g->ctxs[0]->device);
        cudaStreamSynchronize(g->ctxs[0]->stream);
    }
    for (void *p : g->mapped)
    {
        cudaIpcCloseMemHandle(p);
    }
    for (cudaEvent_t e : g->ev_ready)
    {
        if (e)
        {
            cudaEventDestroy(e);
        }
    }
    for (cudaEvent_t e : g->ev_done)
    {
        if (e)
        {
            cudaEventDestroy(e);
        }
    }
    delete g;
    return SPK_OK;
}
} // extern "C"
