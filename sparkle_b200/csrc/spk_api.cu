/*
 * spk_api.cu — the C ABI of include/spk_cuda.h: context lifetime, the device scene builder, camera, accumulation,
 * ray-batch entry points. Every entry point cites the reference interface it stands for in include/spk_cuda.h.
 * No CPU fallback: without a CUDA device spk_cuda_create fails (SPK_ERR_NO_DEVICE).
 */
#include "spk_context.h"

#include <cfloat>
#include <chrono>
#include <cstddef>
#include <cmath>
#include <cstdio>
#include <cstring>
#include <atomic>
#include <memory>
#include <new>
#include <thread>

namespace spk
{
static thread_local std::string g_last_error;

void set_error(const std::string &msg)
{
    g_last_error = msg;
}

int cuda_fail(cudaError_t e, const char *what)
{
    set_error(std::string(what) + ": " + cudaGetErrorString(e));
    return e == cudaErrorMemoryAllocation ? SPK_ERR_OUT_OF_MEMORY : SPK_ERR_CUDA;
}

int ensure_accum_cleared(SpkContext *ctx)
{
    if (ctx->accum_clear_pending)
    {
        SPK_CUDA_TRY(cudaSetDevice(ctx->device));
        SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_accum, 0, size_t(ctx->width) * ctx->height * sizeof(float4), ctx->stream));
        ctx->accum_clear_pending = false;
    }
    return SPK_OK;
}

namespace
{
int invalid(const char *msg)
{
    set_error(msg);
    return SPK_ERR_INVALID_ARGUMENT;
}

struct HostNode
{
    float bounds[6];
    uint32_t index;
};

HostNode read_node(const SpkBvhDesc &d, uint32_t i)
{
    HostNode n;
    memcpy(&n, static_cast<const uint8_t *>(d.nodes) + size_t(i) * d.node_stride, 28);
    return n;
}

uint32_t read_prim_id(const SpkBvhDesc &d, uint32_t k)
{
    return d.prim_id_bytes == 8 ? uint32_t(static_cast<const uint64_t *>(d.prim_ids)[k])
                                : static_cast<const uint32_t *>(d.prim_ids)[k];
}

// structural checks the reference never makes (a malformed tree would overflow its SmallStack<Index,32>)
// host-side flattening of a 1M-triangle scene on one thread was 110 ms of every end-to-end step: work items are spread
// over up to 8 host threads (errors are reported by index, the message is set by the caller's thread)
// (`worth_it` false: a small scene - spawning threads costs more than the work, and with one process per GPU eight of them
// would fight over the same host cores: TestScene's upload went 0.67 -> 0.38 ms)
template <typename Fn> void parallel_for(size_t n, Fn &&fn, bool worth_it = true)
{
    unsigned workers = worth_it ? std::min<size_t>(std::min(8u, std::max(1u, std::thread::hardware_concurrency())), n) : 1u;
    if (workers <= 1)
    {
        for (size_t i = 0; i < n; i++)
        {
            fn(i);
        }
        return;
    }
    std::atomic<size_t> next{0};
    std::vector<std::thread> pool;
    for (unsigned w = 0; w < workers; w++)
    {
        pool.emplace_back([&] {
            for (size_t i = next.fetch_add(1); i < n; i = next.fetch_add(1))
            {
                fn(i);
            }
        });
    }
    for (auto &t : pool)
    {
        t.join();
    }
}

int validate_bvh(const SpkBvhDesc &d, uint32_t expected_prims, const char *what)
{
    if (!d.nodes || d.num_nodes == 0 || (d.node_stride != 28 && d.node_stride != 32) || !d.prim_ids ||
        d.num_prims != expected_prims || (d.prim_id_bytes != 4 && d.prim_id_bytes != 8) || (d.num_nodes & 1u) == 0)
    {
        return invalid(what);
    }
    // iterative depth check
    std::vector<std::pair<uint32_t, uint32_t>> todo;
    todo.emplace_back(0u, 0u);
    uint64_t leaf_prims = 0, visited = 0;
    while (!todo.empty())
    {
        auto [id, depth] = todo.back();
        todo.pop_back();
        if (++visited > d.num_nodes)
        {
            return invalid(what); // children shared between parents (a DAG): a tree visits every node once
        }
        HostNode n = read_node(d, id);
        uint32_t count = n.index & 15u, first = n.index >> 4;
        if (count == 0)
        {
            if ((first & 1u) == 0 || first + 1 >= d.num_nodes || first <= id || depth + 1 > 32)
            {
                return invalid(what);
            }
            todo.emplace_back(first, depth + 1);
            todo.emplace_back(first + 1, depth + 1);
        }
        else
        {
            if (uint64_t(first) + count > d.num_prims)
            {
                return invalid(what);
            }
            leaf_prims += count;
        }
    }
    if (leaf_prims != d.num_prims)
    {
        return invalid(what);
    }
    for (uint32_t k = 0; k < d.num_prims; k++)
    {
        if (read_prim_id(d, k) >= expected_prims)
        {
            return invalid(what);
        }
    }
    return SPK_OK;
}

// nodes[1..] -> 64-byte pairs appended at pair slot `pair_base`; leaf words offset by `leaf_base`.
// By default a pair keeps the builder's depth-first slot. level_order_top (set while SPK_OPT_TRACE_SMEM_TOP is on): slots are
// handed out LEVEL BY LEVEL for the first kTopLevels levels, so the top of the tree is one contiguous run that the ray-batch
// kernel can stage in shared memory (measured and rejected: DESIGN.md 4.1). Index words are absolute, so the visit order of a
// ray does not depend on where a pair is stored.
constexpr uint32_t kTopLevels = 10;

void flatten_pairs(const SpkBvhDesc &d, uint32_t pair_base, uint32_t leaf_base, float *dst, bool level_order_top = false)
{
    const uint32_t num_pairs = (d.num_nodes - 1) / 2;
    std::vector<uint8_t> level(num_pairs, uint8_t(kTopLevels));
    if (level_order_top)
    {
        // level of a pair = depth of its parent node (the pair of the root's children is level 0)
        std::vector<std::pair<uint32_t, uint32_t>> todo;
        todo.emplace_back(0u, 0u);
        while (!todo.empty())
        {
            auto [id, depth] = todo.back();
            todo.pop_back();
            const uint32_t index = read_node(d, id).index;
            if ((index & 15u) == 0u)
            {
                const uint32_t first = index >> 4;
                level[(first - 1) / 2] = uint8_t(std::min(depth, kTopLevels));
                if (depth + 1 < kTopLevels)
                {
                    todo.emplace_back(first, depth + 1);
                    todo.emplace_back(first + 1, depth + 1);
                }
            }
        }
    }
    uint32_t cursor[kTopLevels + 2] = {};
    for (uint32_t q = 0; q < num_pairs; q++)
    {
        cursor[level[q] + 1]++;
    }
    for (uint32_t l = 0; l <= kTopLevels; l++)
    {
        cursor[l + 1] += cursor[l];
    }
    std::vector<uint32_t> slot_of(num_pairs);
    for (uint32_t q = 0; q < num_pairs; q++)
    {
        slot_of[q] = cursor[level[q]]++;
    }
    auto rewrite = [&](uint32_t index) {
        uint32_t count = index & 15u, first = index >> 4;
        return count == 0 ? ((pair_base + slot_of[(first - 1) / 2]) << 4) : (((leaf_base + first) << 4) | count);
    };
    for (uint32_t i = 1; i < d.num_nodes; i++)
    {
        HostNode n = read_node(d, i);
        float *o = dst + (size_t(slot_of[(i - 1) / 2]) * 2 + ((i - 1) & 1u)) * 8;
        memcpy(o, n.bounds, 24);
        uint32_t w = rewrite(n.index);
        memcpy(o + 6, &w, 4);
        o[7] = 0.f;
    }
}

uint32_t root_word(const SpkBvhDesc &d, uint32_t pair_base, uint32_t leaf_base)
{
    HostNode n = read_node(d, 0);
    uint32_t count = n.index & 15u, first = n.index >> 4;
    return count == 0 ? ((pair_base + (first - 1) / 2) << 4) : (((leaf_base + first) << 4) | count);
}

// SpkBvhDesc with nodes == NULL: the tree is built on the device (spk_build.cu)
bool build_on_device(const SpkBvhDesc &d)
{
    return d.nodes == nullptr && d.num_nodes == 0;
}

// pair slots a mesh takes in the one pair array: what its host tree has, or the upper bound of a device build
uint32_t mesh_pair_slots(const SpkMeshDesc &md)
{
    return build_on_device(md.blas) ? md.num_faces : (md.blas.num_nodes - 1) / 2; // n - 1 pairs with one triangle per leaf
}

// ---- scene arena ----
// Every table of a scene lives in ONE device allocation (ctx->d_arena) filled from ONE pinned staging buffer with ONE
// cudaMemcpyAsync and one synchronize. Both are kept across uploads and only grow: re-uploading a scene of the same
// size (the per-frame e2e path, CudaRenderer::Update after a topology change) costs no cudaMalloc / cudaFree at all.
// Round 1 did a cudaMalloc + pageable copy + stream synchronize per table (about 70 per scene) and freed them all on
// the next upload: a fixed 1-3 ms per end-to-end step that did not shrink with the number of GPUs.
struct ArenaPlan
{
    size_t used = 0;
    size_t add(size_t bytes)
    {
        const size_t at = used;
        used += (bytes + 255u) & ~size_t(255u); // 256-byte slots: every table is aligned for 128/256-bit loads
        return at;
    }
};

int arena_reserve(SpkContext *ctx, size_t bytes)
{
    if (ctx->arena_capacity < bytes)
    {
        cudaFree(ctx->d_arena);
        ctx->d_arena = nullptr;
        ctx->arena_capacity = 0;
        const size_t want = bytes + bytes / 8; // head-room for a slightly larger re-upload
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_arena, want));
        ctx->arena_capacity = want;
    }
    if (ctx->stage_capacity < bytes)
    {
        cudaFreeHost(ctx->h_stage);
        ctx->h_stage = nullptr;
        ctx->stage_capacity = 0;
        const size_t want = bytes + bytes / 8;
        SPK_CUDA_TRY(cudaHostAlloc(&ctx->h_stage, want, cudaHostAllocDefault));
        ctx->stage_capacity = want;
    }
    return SPK_OK;
}

size_t texel_bytes(uint32_t format)
{
    switch (format)
    {
    case SPK_FMT_RGBA8_UNORM:
    case SPK_FMT_RGBA8_SRGB:
        return 4;
    case SPK_FMT_RGBA16F:
        return 8;
    case SPK_FMT_RGBA32F:
        return 16;
    default:
        return 0;
    }
}

void free_scene(SpkContext *ctx)
{
    ctx->has_scene = false; // the arena and its staging buffer stay for the next upload (spk_cuda_destroy frees them)
    ctx->scene = DScene{};
    ctx->mesh_root_word.clear();
    ctx->d_pairs = nullptr;
    ctx->d_instances = nullptr;
}

void fill_instance(DInstance &di, const SpkPrimitiveDesc &pd, uint32_t prim, const std::vector<uint32_t> &mesh_root,
                   const std::vector<uint32_t> &material_class)
{
    memset(&di, 0, sizeof(di));
    memcpy(di.inv, pd.inv, 48);
    memcpy(di.fwd, pd.fwd, 48);
    di.kind = pd.kind;
    di.radius2 = pd.radius2;
    di.prim = prim;
    di.material = pd.material;
    di.mesh = pd.kind == SPK_PRIM_MESH ? pd.mesh : SPK_INVALID_ID;
    di.root = pd.kind == SPK_PRIM_MESH ? mesh_root[pd.mesh] : 0u;
    // shade_class (spk_shade.cuh): 1 + mesh + 2 dielectric + 4 textured + 8 emissive, precomputed so the shading sort needs
    // one load per segment instead of instance -> material
    di.shade_class = 1u + (pd.kind == SPK_PRIM_MESH ? 1u : 0u) +
                     (pd.material < material_class.size() ? material_class[pd.material] : 0u);
}
} // namespace
} // namespace spk

using namespace spk;

extern "C"
{
const char *spk_cuda_last_error(void)
{
    return g_last_error.c_str();
}

uint32_t spk_cuda_abi_version(void)
{
    return SPK_ABI_VERSION;
}

int spk_cuda_create(const SpkCreateInfo *info, SpkContext **out)
{
    if (!info || !out || info->abi_version != SPK_ABI_VERSION || info->width < 2 || info->height < 2)
    {
        return invalid("spk_cuda_create: bad create info (abi version, width/height >= 2)");
    }
    *out = nullptr;
    int count = 0;
    cudaError_t e = cudaGetDeviceCount(&count);
    if (e != cudaSuccess || count == 0 || info->device < 0 || info->device >= count)
    {
        set_error("spk_cuda_create: no CUDA device (the cuda pipeline has no CPU fallback)");
        return SPK_ERR_NO_DEVICE;
    }
    SPK_CUDA_TRY(cudaSetDevice(info->device));
    auto *ctx = new (std::nothrow) SpkContext;
    if (!ctx)
    {
        return SPK_ERR_OUT_OF_MEMORY;
    }
    ctx->device = info->device;
    ctx->width = info->width;
    ctx->height = info->height;
    // any failure below releases what was created so far (spk_cuda_destroy accepts a half-built context)
    auto init = [&]() -> int {
        SPK_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
        SPK_CUDA_TRY(cudaEventCreate(&ctx->ev0));
        SPK_CUDA_TRY(cudaEventCreate(&ctx->ev1));
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_accum, size_t(info->width) * info->height * sizeof(float4)));
        SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_accum, 0, size_t(info->width) * info->height * sizeof(float4), ctx->stream));
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_counters, 8 * sizeof(unsigned long long)));
        SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
        SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
        return SPK_OK;
    };
    const int rc = init();
    if (rc != SPK_OK)
    {
        const std::string why = g_last_error;
        spk_cuda_destroy(ctx);
        set_error(why);
        return rc;
    }
    *out = ctx;
    return SPK_OK;
}

int spk_cuda_destroy(SpkContext *ctx)
{
    if (!ctx)
    {
        return SPK_OK;
    }
    cudaSetDevice(ctx->device);
    if (ctx->stream)
    {
        cudaStreamSynchronize(ctx->stream);
    }
    free_scene(ctx);
    cudaFree(ctx->d_arena);
    cudaFreeHost(ctx->h_stage);
    cudaFreeHost(ctx->h_resolve);
    for (cudaEvent_t e : ctx->resolve_ev)
    {
        if (e)
        {
            cudaEventDestroy(e);
        }
    }
    cudaFree(ctx->d_accum);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->d_rays);
    cudaFree(ctx->d_hits);
    if (ctx->copy_in_stream)
    {
        cudaStreamDestroy(ctx->copy_in_stream);
        cudaStreamDestroy(ctx->copy_out_stream);
        for (cudaEvent_t e : ctx->pipe_ev)
        {
            if (e)
            {
                cudaEventDestroy(e);
            }
        }
    }
    cudaFree(ctx->d_resolve);
    cudaFree(ctx->wf_block);
    cudaFree(ctx->d_group_block);
    cudaFree(ctx->d_build_scratch);
    cudaFree(ctx->d_strip);
    cudaFree(ctx->d_gbuf);
    if (ctx->ev0)
    {
        cudaEventDestroy(ctx->ev0);
    }
    if (ctx->ev1)
    {
        cudaEventDestroy(ctx->ev1);
    }
    if (ctx->own_stream && ctx->stream)
    {
        cudaStreamDestroy(ctx->stream);
    }
    delete ctx;
    return SPK_OK;
}

int spk_cuda_set_option(SpkContext *ctx, uint32_t option, int64_t value)
{
    if (!ctx)
    {
        return invalid("null context");
    }
    switch (option)
    {
    case SPK_OPT_MEGAKERNEL_FORM:
        if (value < 0 || value > 9999)
        {
            return invalid("spk_cuda_set_option: megakernel form");
        }
        ctx->opt_mega_form = int(value);
        return SPK_OK;
    case SPK_OPT_WAVEFRONT_SLOTS:
        if (value < 32 || value > (int64_t(1) << 28))
        {
            return invalid("spk_cuda_set_option: wavefront slots must be in [32, 2^28]");
        }
        ctx->opt_wf_slots = uint32_t(value);
        return SPK_OK;
    case SPK_OPT_WAVEFRONT_CHUNK:
        if (value < 1 || value > 1024)
        {
            return invalid("spk_cuda_set_option: wavefront chunk must be in [1, 1024]");
        }
        ctx->opt_wf_chunk = uint32_t(value);
        return SPK_OK;
    case SPK_OPT_WAVEFRONT_SORT:
        if (value < 0 || value > 9)
        {
            return invalid("spk_cuda_set_option: wavefront sort key bits per axis must be in [0, 9]");
        }
        ctx->opt_wf_sort = uint32_t(value);
        return SPK_OK;
    case SPK_OPT_TRACE_VOTE:
        if (value != 4 && value != 6 && value != 8 && value != 12 && value != 16)
        {
            return invalid("spk_cuda_set_option: vote must be 4, 6, 8, 12 or 16");
        }
        ctx->opt_vote = int(value);
        return SPK_OK;
    case SPK_OPT_SMEM_CARVEOUT:
        if (value < -2 || value > 100)
        {
            return invalid("spk_cuda_set_option: carve-out percent");
        }
        ctx->opt_carveout = int(value);
        return SPK_OK;
    case SPK_OPT_VERBOSE:
        ctx->opt_verbose = value != 0;
        return SPK_OK;
    case SPK_OPT_TRACE_SMEM_TOP:
        if (value != 0 && value != 64 && value != 128 && value != 256 && value != 512)
        {
            return invalid("spk_cuda_set_option: staged pairs must be 0, 64, 128, 256 or 512");
        }
        ctx->opt_trace_top = int(value);
        return SPK_OK;
    case SPK_OPT_COST_ORDERED_STRIPS:
        ctx->opt_cost_order = value != 0;
        ctx->strip_order_valid = false;
        return SPK_OK;
    case SPK_OPT_BUILD_MAX_LEAF:
        if (value < 1 || value > 15)
        {
            return invalid("spk_cuda_set_option: leaf size of device-built BLAS must be in [1, 15]");
        }
        ctx->opt_build_max_leaf = uint32_t(value);
        return SPK_OK;
    default:
        return invalid("spk_cuda_set_option: unknown option");
    }
}

int spk_cuda_host_alloc(void **out, size_t bytes)
{
    if (!out || bytes == 0)
    {
        return invalid("spk_cuda_host_alloc: null argument or zero size");
    }
    *out = nullptr;
    SPK_CUDA_TRY(cudaHostAlloc(out, bytes, cudaHostAllocPortable));
    return SPK_OK;
}

int spk_cuda_host_free(void *p)
{
    if (p)
    {
        SPK_CUDA_TRY(cudaFreeHost(p));
    }
    return SPK_OK;
}

int spk_cuda_set_stream(SpkContext *ctx, void *cuda_stream)
{
    if (!ctx)
    {
        return invalid("null context");
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    if (ctx->own_stream)
    {
        cudaStreamDestroy(ctx->stream);
    }
    ctx->stream = static_cast<cudaStream_t>(cuda_stream);
    ctx->own_stream = false;
    return SPK_OK;
}
} // extern "C"

namespace spk
{
// The run of pair slots the ray-batch kernel may stage in shared memory (SPK_OPT_TRACE_SMEM_TOP): the TLAS when the scene
// has one with inner nodes (every ray starts there), else the top levels of the first mesh's BLAS.
void choose_top_region(SpkContext *ctx)
{
    ctx->scene.top_base = 0;
    ctx->scene.top_count = 0;
    if (ctx->tlas_tree.num_pairs > 0)
    {
        ctx->scene.top_base = ctx->tlas_tree.pair_base;
        ctx->scene.top_count = ctx->tlas_tree.num_pairs;
    }
    else if (!ctx->mesh_trees.empty() && ctx->mesh_trees[0].num_pairs > 0)
    {
        ctx->scene.top_base = ctx->mesh_trees[0].pair_base;
        ctx->scene.top_count = ctx->mesh_trees[0].num_pairs;
    }
}

// TLAS::Build on the device: world boxes of the primitives (AABB::TransformTo of the mesh bounds, SceneRenderProxy.cpp:41-47;
// centre +- radius for spheres) -> LBVH with one primitive per leaf -> instances in the order of the sorted leaves.
int rebuild_tlas_on_device(SpkContext *ctx, const SpkPrimitiveDesc *prims, uint32_t n)
{
    std::vector<float> boxes(size_t(n) * 6);
    for (uint32_t p = 0; p < n; p++)
    {
        const SpkPrimitiveDesc &pd = prims[p];
        float lo[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, hi[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        if (pd.kind == SPK_PRIM_SPHERE)
        {
            const float r = std::sqrt(pd.radius2) * 1.000001f; // never inside the sphere by a rounding of the square root
            for (int a = 0; a < 3; a++)
            {
                lo[a] = pd.fwd[4 * a + 3] - r;
                hi[a] = pd.fwd[4 * a + 3] + r;
            }
        }
        else
        {
            const float *b = ctx->mesh_trees[pd.mesh].root_bounds;
            for (int corner = 0; corner < 8; corner++)
            {
                const float c[3] = {corner & 4 ? b[3] : b[0], corner & 2 ? b[4] : b[1], corner & 1 ? b[5] : b[2]};
                for (int a = 0; a < 3; a++)
                {
                    const float *row = pd.fwd + 4 * a;
                    const float w = ((row[0] * c[0] + row[1] * c[1]) + row[2] * c[2]) + row[3];
                    lo[a] = std::min(lo[a], w);
                    hi[a] = std::max(hi[a], w);
                }
            }
        }
        memcpy(&boxes[size_t(p) * 6], lo, 12);
        memcpy(&boxes[size_t(p) * 6 + 3], hi, 12);
    }
    SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_tlas_boxes, boxes.data(), boxes.size() * 4, cudaMemcpyHostToDevice, ctx->stream));
    DeviceBuildResult res;
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev0, ctx->stream));
    int rc = device_build_lbvh(ctx, nullptr, nullptr, ctx->d_tlas_boxes, n, 1u, ctx->d_pairs, ctx->tlas_tree.pair_base, 0u,
                               nullptr, &res);
    if (rc != SPK_OK)
    {
        return rc;
    }
    SPK_CUDA_TRY(cudaEventRecord(ctx->ev1, ctx->stream));
    std::vector<uint32_t> order(n);
    SPK_CUDA_TRY(cudaMemcpyAsync(order.data(), ctx->build_ids, size_t(n) * 4, cudaMemcpyDeviceToHost, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    float ms = 0.f;
    cudaEventElapsedTime(&ms, ctx->ev0, ctx->ev1);
    std::vector<DInstance> instances(n);
    for (uint32_t k = 0; k < n; k++)
    {
        fill_instance(instances[k], prims[order[k]], order[k], ctx->mesh_root_word, ctx->material_class);
    }
    SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_instances, instances.data(), instances.size() * sizeof(DInstance),
                                 cudaMemcpyHostToDevice, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    SpkContext::TreeInfo &t = ctx->tlas_tree;
    t.num_pairs = res.num_pairs;
    t.root_word = res.root_word;
    t.depth = res.depth;
    t.sah_cost = res.sah_cost;
    t.build_ms = ms;
    t.device_built = true;
    t.num_prims = n;
    memcpy(t.root_bounds, res.root_bounds, 24);
    ctx->scene.tlas_root = res.root_word;
    return SPK_OK;
}
} // namespace spk

/* ---- device scene builder ------------------------------------------------------------------------------------- */
extern "C"
{
int spk_cuda_upload_scene(SpkContext *ctx, const SpkSceneDesc *s)
{
    if (!ctx || !s || s->abi_version != SPK_ABI_VERSION)
    {
        return invalid("spk_cuda_upload_scene: null argument or ABI version mismatch");
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    free_scene(ctx);
    int rc;
    // SPK_OPT_VERBOSE: where the time of an upload goes (host clock)
    auto now = [] { return std::chrono::steady_clock::now(); };
    auto ms_since = [&](std::chrono::steady_clock::time_point t) {
        return std::chrono::duration<double, std::milli>(now() - t).count();
    };
    const auto t_begin = now();

    // ---- validation (before any allocation) ----
    if ((s->num_meshes && !s->meshes) || (s->num_primitives && !s->primitives) || (s->num_materials && !s->materials) ||
        (s->num_textures && !s->textures))
    {
        return invalid("spk_cuda_upload_scene: null table");
    }
    uint64_t total_pairs = 0, total_tris = 0;
    std::vector<int> mesh_status(s->num_meshes, 0); // 0 ok, 1 missing arrays, 2 index out of range, 3 malformed BLAS
    uint64_t faces_in = 0;
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        faces_in += s->meshes[m].num_faces;
    }
    const bool big_scene = faces_in > 100000u; // host threads pay off from ~5 MB of tables
    parallel_for(s->num_meshes, [&](size_t m) {
        const SpkMeshDesc &md = s->meshes[m];
        if (!md.vertices || !md.normals || !md.tangents || !md.indices || md.num_faces == 0 || md.num_vertices == 0)
        {
            mesh_status[m] = 1;
            return;
        }
        for (size_t k = 0; k < size_t(md.num_faces) * 3; k++)
        {
            if (md.indices[k] >= md.num_vertices)
            {
                mesh_status[m] = 2;
                return;
            }
        }
        if (build_on_device(md.blas))
        {
            return; // BLAS::Build runs on the device (spk_build.cu)
        }
        if (validate_bvh(md.blas, md.num_faces, "") != SPK_OK)
        {
            mesh_status[m] = 3;
        }
    }, big_scene);
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        if (mesh_status[m] != 0)
        {
            return invalid(mesh_status[m] == 1   ? "spk_cuda_upload_scene: mesh needs vertices, normals, tangents, indices"
                           : mesh_status[m] == 2 ? "spk_cuda_upload_scene: vertex index out of range"
                                                 : "spk_cuda_upload_scene: malformed BLAS");
        }
        total_pairs += mesh_pair_slots(s->meshes[m]);
        total_tris += s->meshes[m].num_faces;
    }
    for (uint32_t p = 0; p < s->num_primitives; p++)
    {
        const SpkPrimitiveDesc &pd = s->primitives[p];
        if (pd.kind > SPK_PRIM_SPHERE || pd.material >= s->num_materials ||
            (pd.kind == SPK_PRIM_MESH && pd.mesh >= s->num_meshes))
        {
            return invalid("spk_cuda_upload_scene: primitive references a missing mesh or material");
        }
    }
    for (uint32_t m = 0; m < s->num_materials; m++)
    {
        const SpkMaterialDesc &md = s->materials[m];
        const uint32_t ids[4] = {md.base_color_tex, md.emissive_tex, md.metallic_roughness_tex, md.normal_tex};
        for (uint32_t id : ids)
        {
            if (id != SPK_INVALID_ID && id >= s->num_textures)
            {
                return invalid("spk_cuda_upload_scene: material references a missing texture");
            }
        }
        if (md.type > SPK_MAT_DIELECTRIC)
        {
            return invalid("spk_cuda_upload_scene: material type");
        }
    }
    const bool tlas_on_device = s->num_primitives && build_on_device(s->tlas);
    if (s->num_primitives)
    {
        if (tlas_on_device)
        {
            total_pairs += s->num_primitives; // upper bound of a rebuilt TLAS (n - 1 pairs)
        }
        else
        {
            if ((rc = validate_bvh(s->tlas, s->num_primitives, "spk_cuda_upload_scene: malformed TLAS")) != SPK_OK)
            {
                return rc;
            }
            // a later spk_cuda_update_instances may rebuild the TLAS on the device: leave room for its n - 1 pairs
            total_pairs += std::max<uint32_t>((s->tlas.num_nodes - 1) / 2, s->num_primitives);
        }
    }
    if (total_pairs >= (1ull << 28) || total_tris >= (1ull << 28))
    {
        return invalid("spk_cuda_upload_scene: scene exceeds 2^28 pairs or triangles");
    }

    // ---- texture descriptors are checked before anything is laid out ----
    auto texture_ok = [](const SpkTextureDesc &t) { return t.pixels && t.width && t.height && texel_bytes(t.format); };
    for (uint32_t t = 0; t < s->num_textures; t++)
    {
        if (!texture_ok(s->textures[t]))
        {
            return invalid("texture descriptor");
        }
    }
    const bool sky_map = s->sky.has_sky && s->sky.has_map;
    for (int f = 0; sky_map && f < 6; f++)
    {
        if (!texture_ok(s->sky.faces[f]))
        {
            return invalid("texture descriptor");
        }
    }

    const double ms_validate = ms_since(t_begin);
    const auto t_fill = now();
    // ---- lay the arena out (offsets only) ----
    ArenaPlan plan;
    // one spare pair so an all-leaf scene still has a valid (never dereferenced) pointer
    const size_t off_pairs = plan.add(size_t(total_pairs) * 64 + 64);
    const size_t off_tris = plan.add(size_t(total_tris) * 48);
    const size_t off_inst = plan.add(size_t(s->num_primitives) * sizeof(DInstance));
    const size_t off_meshes = plan.add(size_t(s->num_meshes) * sizeof(DMesh));
    const size_t off_mats = plan.add(size_t(s->num_materials) * sizeof(DMaterial));
    const size_t off_tex = plan.add((size_t(s->num_textures) + 6) * sizeof(DTexture)); // + the six faces of a sky map
    struct MeshOffsets
    {
        size_t normals, tangents, indices, uvs, vertices;
    };
    std::vector<MeshOffsets> mesh_off(s->num_meshes);
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        const SpkMeshDesc &md = s->meshes[m];
        mesh_off[m].normals = plan.add(size_t(md.num_vertices) * 12);
        mesh_off[m].tangents = plan.add(size_t(md.num_vertices) * 16);
        mesh_off[m].indices = plan.add(size_t(md.num_faces) * 12);
        mesh_off[m].uvs = md.uvs ? plan.add(size_t(md.num_vertices) * 8) : 0;
        mesh_off[m].vertices = build_on_device(md.blas) ? plan.add(size_t(md.num_vertices) * 12) : 0;
    }
    const size_t off_boxes = plan.add(size_t(s->num_primitives) * 24); // world boxes of a device-built TLAS
    auto texture_bytes = [](const SpkTextureDesc &t) { return size_t(t.width) * t.height * texel_bytes(t.format); };
    std::vector<size_t> tex_off(s->num_textures);
    for (uint32_t t = 0; t < s->num_textures; t++)
    {
        tex_off[t] = plan.add(texture_bytes(s->textures[t]));
    }
    size_t sky_off[6] = {0, 0, 0, 0, 0, 0};
    for (int f = 0; sky_map && f < 6; f++)
    {
        sky_off[f] = plan.add(texture_bytes(s->sky.faces[f]));
    }
    if ((rc = arena_reserve(ctx, plan.used)) != SPK_OK)
    {
        return rc;
    }
    uint8_t *const stage = static_cast<uint8_t *>(ctx->h_stage);
    uint8_t *const dev = static_cast<uint8_t *>(ctx->d_arena);

    // ---- flatten BLAS pairs + triangles (leaf order) straight into the pinned staging buffer ----
    // (every word of both tables is written below; nothing is zero-filled first)
    float *const pairs = reinterpret_cast<float *>(stage + off_pairs);
    float *const tris = reinterpret_cast<float *>(stage + off_tris);
    memset(pairs + size_t(total_pairs) * 16, 0, 64);
    ctx->mesh_root_word.assign(s->num_meshes, 0);
    ctx->mesh_trees.assign(s->num_meshes, SpkContext::TreeInfo{});
    ctx->tlas_tree = SpkContext::TreeInfo{};
    uint32_t pair_base = 0, tri_base = 0;
    // work items: one flatten_pairs per mesh, the triangle records in chunks of 64 Ki faces, and plain copies of the
    // attribute / texture arrays in chunks of 4 MiB
    struct Item
    {
        uint32_t mesh, pair_base, tri_base, begin, end; // begin == end: the pair item
        const void *src = nullptr;                     // != nullptr: a plain copy of `bytes` to stage + dst
        size_t dst = 0, bytes = 0;
    };
    std::vector<Item> items;
    auto add_copy = [&](const void *src, size_t dst, size_t bytes) {
        const size_t chunk = size_t(4) << 20;
        for (size_t at = 0; at < bytes; at += chunk)
        {
            Item it{};
            it.src = static_cast<const uint8_t *>(src) + at;
            it.dst = dst + at;
            it.bytes = std::min(chunk, bytes - at);
            items.push_back(it);
        }
    };
    DMesh *const dmeshes = reinterpret_cast<DMesh *>(stage + off_meshes);
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        const SpkMeshDesc &md = s->meshes[m];
        SpkContext::TreeInfo &tree = ctx->mesh_trees[m];
        tree.pair_base = pair_base;
        tree.leaf_base = tri_base;
        tree.num_prims = md.num_faces;
        tree.device_built = build_on_device(md.blas);
        if (tree.device_built)
        {
            add_copy(md.vertices, mesh_off[m].vertices, size_t(md.num_vertices) * 12);
        }
        else
        {
            items.push_back(Item{m, pair_base, tri_base, 0, 0});
            for (uint32_t b = 0; b < md.num_faces; b += 65536u)
            {
                items.push_back(Item{m, pair_base, tri_base, b, std::min(md.num_faces, b + 65536u)});
            }
            tree.num_pairs = (md.blas.num_nodes - 1) / 2;
            const HostNode root = read_node(md.blas, 0);
            for (int a = 0; a < 3; a++)
            {
                tree.root_bounds[a] = root.bounds[2 * a];
                tree.root_bounds[3 + a] = root.bounds[2 * a + 1];
            }
        }
        pair_base += mesh_pair_slots(md);
        tri_base += md.num_faces;
        add_copy(md.normals, mesh_off[m].normals, size_t(md.num_vertices) * 12);
        add_copy(md.tangents, mesh_off[m].tangents, size_t(md.num_vertices) * 16);
        add_copy(md.indices, mesh_off[m].indices, size_t(md.num_faces) * 12);
        if (md.uvs)
        {
            add_copy(md.uvs, mesh_off[m].uvs, size_t(md.num_vertices) * 8);
        }
        dmeshes[m].normals = reinterpret_cast<const float *>(dev + mesh_off[m].normals);
        dmeshes[m].tangents = reinterpret_cast<const float *>(dev + mesh_off[m].tangents);
        dmeshes[m].indices = reinterpret_cast<const uint32_t *>(dev + mesh_off[m].indices);
        dmeshes[m].uvs = md.uvs ? reinterpret_cast<const float *>(dev + mesh_off[m].uvs) : nullptr;
    }
    DTexture *const dtex = reinterpret_cast<DTexture *>(stage + off_tex);
    auto device_texture = [&](const SpkTextureDesc &t, size_t off) {
        DTexture d;
        d.pixels = dev + off;
        d.width = t.width;
        d.height = t.height;
        d.format = t.format;
        d.pad = 0;
        return d;
    };
    for (uint32_t t = 0; t < s->num_textures; t++)
    {
        add_copy(s->textures[t].pixels, tex_off[t], texture_bytes(s->textures[t]));
        dtex[t] = device_texture(s->textures[t], tex_off[t]);
    }
    DSky sky{};
    sky.has_sky = s->sky.has_sky;
    sky.has_map = sky_map ? 1u : 0u;
    memcpy(sky.color, s->sky.color, 12);
    for (int f = 0; sky_map && f < 6; f++)
    {
        add_copy(s->sky.faces[f].pixels, sky_off[f], texture_bytes(s->sky.faces[f]));
        dtex[s->num_textures + f] = device_texture(s->sky.faces[f], sky_off[f]);
    }
    sky.face_base = s->num_textures;
    static_assert(sizeof(SpkMaterialDesc) == sizeof(DMaterial), "material ABI == device record");
    if (s->num_materials)
    {
        memcpy(stage + off_mats, s->materials, size_t(s->num_materials) * sizeof(DMaterial));
    }
    parallel_for(items.size(), [&](size_t w) {
        const Item &it = items[w];
        if (it.src)
        {
            memcpy(stage + it.dst, it.src, it.bytes);
            return;
        }
        const SpkMeshDesc &md = s->meshes[it.mesh];
        if (it.begin == it.end)
        {
            flatten_pairs(md.blas, it.pair_base, it.tri_base, pairs + size_t(it.pair_base) * 16, ctx->opt_trace_top > 0);
            ctx->mesh_root_word[it.mesh] = root_word(md.blas, it.pair_base, it.tri_base);
            ctx->mesh_trees[it.mesh].root_word = ctx->mesh_root_word[it.mesh];
            return;
        }
        for (uint32_t k = it.begin; k < it.end; k++)
        {
            uint32_t face = read_prim_id(md.blas, k);
            const float *v0 = md.vertices + size_t(md.indices[size_t(face) * 3 + 0]) * 3;
            const float *v1 = md.vertices + size_t(md.indices[size_t(face) * 3 + 1]) * 3;
            const float *v2 = md.vertices + size_t(md.indices[size_t(face) * 3 + 2]) * 3;
            float *o = tris + size_t(it.tri_base + k) * 12;
            // Triangle::Set (MeshRenderProxy.cpp:80-86): p0 = v0, e1 = v1 - v0, e2 = v2 - v0
            for (int a = 0; a < 3; a++)
            {
                o[a] = v0[a];
                o[4 + a] = v1[a] - v0[a];
                o[8 + a] = v2[a] - v0[a];
            }
            memcpy(o + 3, &face, 4);
            o[7] = 0.f;
            o[11] = 0.f;
        }
    }, big_scene || plan.used > (size_t(8) << 20));
    ctx->num_blas_pairs = pair_base;
    ctx->tlas_pair_base = pair_base;

    ctx->material_class.assign(s->num_materials, 0u);
    for (uint32_t m = 0; m < s->num_materials; m++)
    {
        const SpkMaterialDesc &md = s->materials[m];
        const bool textured = (md.base_color_tex & md.metallic_roughness_tex & md.normal_tex & md.emissive_tex) != SPK_INVALID_ID;
        const bool emissive = md.emissive[0] != 0.f || md.emissive[1] != 0.f || md.emissive[2] != 0.f;
        ctx->material_class[m] = (md.type == SPK_MAT_DIELECTRIC ? 2u : 0u) + (textured ? 4u : 0u) + (emissive ? 8u : 0u);
    }
    // ---- TLAS pairs + instances in TLAS leaf order (SceneRenderProxy.cpp:59-68) ----
    DInstance *const instances = reinterpret_cast<DInstance *>(stage + off_inst);
    ctx->scene = DScene{};
    ctx->tlas_tree.pair_base = pair_base;
    ctx->tlas_tree.num_prims = s->num_primitives;
    bool meshes_on_device = false;
    for (const SpkContext::TreeInfo &t : ctx->mesh_trees)
    {
        meshes_on_device |= t.device_built;
    }
    if (s->num_primitives && !tlas_on_device)
    {
        flatten_pairs(s->tlas, pair_base, 0, pairs + size_t(pair_base) * 16);
        ctx->scene.tlas_root = root_word(s->tlas, pair_base, 0);
        ctx->tlas_tree.root_word = ctx->scene.tlas_root;
        ctx->tlas_tree.num_pairs = (s->tlas.num_nodes - 1) / 2;
        const HostNode troot = read_node(s->tlas, 0);
        for (int a = 0; a < 3; a++)
        {
            ctx->tlas_tree.root_bounds[a] = troot.bounds[2 * a];
            ctx->tlas_tree.root_bounds[3 + a] = troot.bounds[2 * a + 1];
        }
        if (!meshes_on_device)
        {
            for (uint32_t k = 0; k < s->num_primitives; k++)
            {
                uint32_t prim = read_prim_id(s->tlas, k);
                fill_instance(instances[k], s->primitives[prim], prim, ctx->mesh_root_word, ctx->material_class);
            }
        }
    }
    ctx->pair_capacity = uint32_t(total_pairs);
    ctx->instance_capacity = s->num_primitives;

    const double ms_fill = ms_since(t_fill);
    const auto t_copy = now();
    // ---- one copy, one synchronize (host pointers are borrowed for this call only) ----
    SPK_CUDA_TRY(cudaMemcpyAsync(dev, stage, plan.used, cudaMemcpyHostToDevice, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    if (ctx->opt_verbose)
    {
        fprintf(stderr, "[spk_cuda] upload_scene: validate %.2f ms, flatten + stage %.2f ms, copy %.1f MB %.2f ms\n", ms_validate,
                ms_fill, plan.used / 1e6, ms_since(t_copy));
    }
    ctx->d_pairs = reinterpret_cast<float4 *>(dev + off_pairs);
    ctx->d_instances = reinterpret_cast<DInstance *>(dev + off_inst);
    ctx->d_tlas_boxes = reinterpret_cast<float *>(dev + off_boxes);

    // ---- BLAS::Build on the device for the meshes that came without a tree (MeshRenderProxy.cpp:31-65) ----
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        SpkContext::TreeInfo &tree = ctx->mesh_trees[m];
        if (!tree.device_built)
        {
            continue;
        }
        DeviceBuildResult res;
        cudaEvent_t e0 = ctx->ev0, e1 = ctx->ev1;
        SPK_CUDA_TRY(cudaEventRecord(e0, ctx->stream));
        // one triangle per leaf traces fastest (tools/r2_lbvh_leaf_ab.py) but makes the deepest tree: a tree that does not
        // fit the 32-entry traversal stack is rebuilt with larger leaves (each doubling takes about one level off)
        for (uint32_t leaf = ctx->opt_build_max_leaf;; leaf = std::min(15u, leaf * 2u))
        {
            rc = device_build_lbvh(ctx, reinterpret_cast<const float *>(dev + mesh_off[m].vertices),
                                   reinterpret_cast<const uint32_t *>(dev + mesh_off[m].indices), nullptr,
                                   s->meshes[m].num_faces, leaf, ctx->d_pairs, tree.pair_base, tree.leaf_base,
                                   reinterpret_cast<float4 *>(dev + off_tris) + size_t(tree.leaf_base) * 3, &res);
            if (rc != SPK_ERR_UNSUPPORTED || leaf >= 15u)
            {
                break;
            }
        }
        if (rc != SPK_OK)
        {
            return rc;
        }
        SPK_CUDA_TRY(cudaEventRecord(e1, ctx->stream));
        SPK_CUDA_TRY(cudaEventSynchronize(e1));
        float ms = 0.f;
        cudaEventElapsedTime(&ms, e0, e1);
        tree.build_ms = ms;
        tree.num_pairs = res.num_pairs;
        tree.root_word = res.root_word;
        tree.depth = res.depth;
        tree.sah_cost = res.sah_cost;
        memcpy(tree.root_bounds, res.root_bounds, 24);
        ctx->mesh_root_word[m] = res.root_word;
    }
    // ---- TLAS::Build on the device (SceneRenderProxy.cpp:31-69), or instances that waited for device-built BLAS roots ----
    if (tlas_on_device)
    {
        if ((rc = rebuild_tlas_on_device(ctx, s->primitives, s->num_primitives)) != SPK_OK)
        {
            return rc;
        }
    }
    else if (s->num_primitives && meshes_on_device)
    {
        std::vector<DInstance> late(s->num_primitives);
        for (uint32_t k = 0; k < s->num_primitives; k++)
        {
            uint32_t prim = read_prim_id(s->tlas, k);
            fill_instance(late[k], s->primitives[prim], prim, ctx->mesh_root_word, ctx->material_class);
        }
        SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_instances, late.data(), late.size() * sizeof(DInstance), cudaMemcpyHostToDevice,
                                     ctx->stream));
        SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    }

    const uint32_t tlas_root_keep = ctx->scene.tlas_root;
    ctx->scene.pairs = ctx->d_pairs;
    ctx->scene.tris = reinterpret_cast<const float4 *>(dev + off_tris);
    ctx->scene.instances = ctx->d_instances;
    ctx->scene.meshes = reinterpret_cast<const DMesh *>(dev + off_meshes);
    ctx->scene.materials = reinterpret_cast<const DMaterial *>(dev + off_mats);
    ctx->scene.textures = reinterpret_cast<const DTexture *>(dev + off_tex);
    ctx->scene.num_instances = s->num_primitives;
    ctx->scene.tlas_root = tlas_root_keep;
    ctx->scene.cull_extent = s->world_extent > 0.f ? s->world_extent : 1.f;
    ctx->scene.sky = sky;
    choose_top_region(ctx);
    ctx->has_scene = true;
    ctx->tracer_choice = -1; // measured again for the new scene (spk_cuda_choose_tracer)
    return SPK_OK;
}

int spk_cuda_update_instances(SpkContext *ctx, const SpkPrimitiveDesc *primitives, uint32_t num_primitives,
                              const SpkBvhDesc *tlas)
{
    if (!ctx || !primitives)
    {
        return invalid("spk_cuda_update_instances: null argument");
    }
    if (!ctx->has_scene)
    {
        set_error("spk_cuda_update_instances: no scene uploaded");
        return SPK_ERR_NO_SCENE;
    }
    if (num_primitives != ctx->instance_capacity || num_primitives == 0)
    {
        return invalid("spk_cuda_update_instances: primitive count changed (use spk_cuda_upload_scene)");
    }
    for (uint32_t p = 0; p < num_primitives; p++)
    {
        if (primitives[p].kind > SPK_PRIM_SPHERE || primitives[p].material >= ctx->material_class.size() ||
            (primitives[p].kind == SPK_PRIM_MESH && primitives[p].mesh >= ctx->mesh_root_word.size()))
        {
            return invalid("spk_cuda_update_instances: primitive references a missing mesh or material");
        }
    }
    if (!tlas || build_on_device(*tlas))
    {
        // TLAS == NULL: rebuilt on the device from the new transforms (SceneRenderProxy.cpp:212-230, 269-297)
        SPK_CUDA_TRY(cudaSetDevice(ctx->device));
        int rc0 = rebuild_tlas_on_device(ctx, primitives, num_primitives);
        if (rc0 != SPK_OK)
        {
            return rc0;
        }
        float extent = 0.f;
        for (float b : ctx->tlas_tree.root_bounds)
        {
            extent = std::max(extent, std::fabs(b));
        }
        if (extent > 0.f && std::isfinite(extent))
        {
            ctx->scene.cull_extent = extent;
        }
        choose_top_region(ctx);
        return SPK_OK;
    }
    int rc = validate_bvh(*tlas, num_primitives, "spk_cuda_update_instances: malformed TLAS");
    if (rc != SPK_OK)
    {
        return rc;
    }
    uint32_t tlas_pairs = (tlas->num_nodes - 1) / 2;
    if (ctx->tlas_pair_base + tlas_pairs > ctx->pair_capacity)
    {
        return invalid("spk_cuda_update_instances: rebuilt TLAS has more nodes than the uploaded one");
    }
    for (uint32_t p = 0; p < num_primitives; p++)
    {
        if (primitives[p].kind > SPK_PRIM_SPHERE || primitives[p].material >= ctx->material_class.size() ||
            (primitives[p].kind == SPK_PRIM_MESH && primitives[p].mesh >= ctx->mesh_root_word.size()))
        {
            return invalid("spk_cuda_update_instances: primitive references a missing mesh or material");
        }
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    std::vector<float> pairs(size_t(tlas_pairs) * 16 + 16, 0.f);
    flatten_pairs(*tlas, ctx->tlas_pair_base, 0, pairs.data());
    std::vector<DInstance> instances(num_primitives);
    for (uint32_t k = 0; k < num_primitives; k++)
    {
        uint32_t prim = read_prim_id(*tlas, k);
        fill_instance(instances[k], primitives[prim], prim, ctx->mesh_root_word, ctx->material_class);
    }
    if (tlas_pairs)
    {
        SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_pairs + size_t(ctx->tlas_pair_base) * 4, pairs.data(),
                                     size_t(tlas_pairs) * 64, cudaMemcpyHostToDevice, ctx->stream));
    }
    SPK_CUDA_TRY(cudaMemcpyAsync(ctx->d_instances, instances.data(), instances.size() * sizeof(DInstance),
                                 cudaMemcpyHostToDevice, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    ctx->scene.tlas_root = root_word(*tlas, ctx->tlas_pair_base, 0);
    ctx->tlas_tree.root_word = ctx->scene.tlas_root;
    ctx->tlas_tree.num_pairs = tlas_pairs;
    ctx->tlas_tree.device_built = false;
    choose_top_region(ctx);
    // the t-culling slack is sized from the world extent (spk_traverse.cuh): primitives may have moved outside the
    // bounds of the upload, so take it again from the rebuilt TLAS root (= the world AABB)
    const HostNode root = read_node(*tlas, 0);
    float extent = 0.f;
    for (float b : root.bounds)
    {
        extent = std::max(extent, std::fabs(b));
    }
    for (int a = 0; a < 3; a++)
    {
        ctx->tlas_tree.root_bounds[a] = root.bounds[2 * a];
        ctx->tlas_tree.root_bounds[3 + a] = root.bounds[2 * a + 1];
    }
    if (extent > 0.f && std::isfinite(extent))
    {
        ctx->scene.cull_extent = extent;
    }
    return SPK_OK;
}

int spk_cuda_download_bvh(SpkContext *ctx, uint32_t mesh, SpkBvhNode *nodes_out, uint32_t node_capacity,
                          uint32_t *num_nodes_out, uint32_t *prim_ids_out, uint32_t prim_capacity, uint32_t *num_prims_out)
{
    if (!ctx || !num_nodes_out || !num_prims_out)
    {
        return invalid("spk_cuda_download_bvh: null argument");
    }
    if (!ctx->has_scene)
    {
        set_error("spk_cuda_download_bvh: no scene uploaded");
        return SPK_ERR_NO_SCENE;
    }
    const bool is_tlas = mesh == SPK_BVH_TLAS;
    if (!is_tlas && mesh >= ctx->mesh_trees.size())
    {
        return invalid("spk_cuda_download_bvh: no such mesh");
    }
    const SpkContext::TreeInfo &t = is_tlas ? ctx->tlas_tree : ctx->mesh_trees[mesh];
    *num_nodes_out = 1 + 2 * t.num_pairs;
    *num_prims_out = t.num_prims;
    if (!nodes_out && !prim_ids_out)
    {
        return SPK_OK; // size query
    }
    if (!nodes_out || !prim_ids_out || node_capacity < *num_nodes_out || prim_capacity < t.num_prims)
    {
        return invalid("spk_cuda_download_bvh: output arrays too small (call with NULL arrays for the sizes)");
    }
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    std::vector<float> pairs(size_t(t.num_pairs) * 16);
    if (t.num_pairs)
    {
        SPK_CUDA_TRY(cudaMemcpyAsync(pairs.data(), ctx->d_pairs + size_t(t.pair_base) * 4, pairs.size() * 4,
                                     cudaMemcpyDeviceToHost, ctx->stream));
    }
    // primitive ids in leaf order: the face index in a triangle record, the primitive index in an instance record
    if (t.num_prims)
    {
        const uint8_t *src = is_tlas ? reinterpret_cast<const uint8_t *>(ctx->d_instances) + offsetof(DInstance, prim)
                                     : reinterpret_cast<const uint8_t *>(ctx->scene.tris) + size_t(t.leaf_base) * 48 + 12;
        SPK_CUDA_TRY(cudaMemcpy2DAsync(prim_ids_out, 4, src, is_tlas ? sizeof(DInstance) : 48, 4, t.num_prims,
                                       cudaMemcpyDeviceToHost, ctx->stream));
    }
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
    // absolute index words back to the bvh::v2 form: children of an inner node at ids 1 + 2 * slot, leaves relative
    auto relative = [&](uint32_t word) {
        const uint32_t count = word & 15u, first = word >> 4;
        return count == 0 ? ((1u + 2u * (first - t.pair_base)) << 4) : (((first - t.leaf_base) << 4) | count);
    };
    memset(nodes_out, 0, sizeof(SpkBvhNode) * size_t(*num_nodes_out));
    for (int a = 0; a < 3; a++)
    {
        nodes_out[0].bounds[2 * a] = t.root_bounds[a];
        nodes_out[0].bounds[2 * a + 1] = t.root_bounds[3 + a];
    }
    nodes_out[0].index = relative(t.root_word);
    for (uint32_t k = 0; k < 2 * t.num_pairs; k++)
    {
        const float *rec = pairs.data() + size_t(k) * 8;
        memcpy(nodes_out[1 + k].bounds, rec, 24);
        uint32_t word;
        memcpy(&word, rec + 6, 4);
        nodes_out[1 + k].index = relative(word);
    }
    return SPK_OK;
}

int spk_cuda_get_build_info(SpkContext *ctx, uint32_t mesh, SpkBuildInfo *out)
{
    if (!ctx || !out)
    {
        return invalid("spk_cuda_get_build_info: null argument");
    }
    if (!ctx->has_scene)
    {
        set_error("spk_cuda_get_build_info: no scene uploaded");
        return SPK_ERR_NO_SCENE;
    }
    const bool is_tlas = mesh == SPK_BVH_TLAS;
    if (!is_tlas && mesh >= ctx->mesh_trees.size())
    {
        return invalid("spk_cuda_get_build_info: no such mesh");
    }
    const SpkContext::TreeInfo &t = is_tlas ? ctx->tlas_tree : ctx->mesh_trees[mesh];
    out->device_built = t.device_built ? 1u : 0u;
    out->num_nodes = 1 + 2 * t.num_pairs;
    out->num_prims = t.num_prims;
    out->depth = t.depth;
    out->sah_cost = t.sah_cost;
    out->build_ms = t.build_ms;
    return SPK_OK;
}

int spk_cuda_set_camera(SpkContext *ctx, const SpkCamera *camera)
{
    if (!ctx || !camera)
    {
        return invalid("spk_cuda_set_camera: null argument");
    }
    static_assert(sizeof(SpkCamera) == sizeof(DCamera), "camera ABI == device record");
    memcpy(&ctx->camera, camera, sizeof(DCamera));
    ctx->has_camera = true;
    return SPK_OK;
}

int spk_cuda_reset_accumulation(SpkContext *ctx)
{
    if (!ctx)
    {
        return invalid("null context");
    }
    ctx->accum_clear_pending = true; // the next megakernel launch starts from zero; other consumers clear on demand
    ctx->accum_samples = 0;
    return SPK_OK;
}

int spk_cuda_accum_device_ptr(SpkContext *ctx, void **device_ptr, size_t *bytes)
{
    if (!ctx || !device_ptr || !bytes)
    {
        return invalid("spk_cuda_accum_device_ptr: null argument");
    }
    int rc = ensure_accum_cleared(ctx); // an external reader (NCCL reduce) sees zeros after a reset
    if (rc != SPK_OK)
    {
        return rc;
    }
    *device_ptr = ctx->d_accum;
    *bytes = size_t(ctx->width) * ctx->height * sizeof(float4);
    return SPK_OK;
}

int spk_cuda_accum_sample_count(SpkContext *ctx, uint32_t *count)
{
    if (!ctx || !count)
    {
        return invalid("spk_cuda_accum_sample_count: null argument");
    }
    *count = ctx->accum_samples;
    return SPK_OK;
}
} // extern "C"
