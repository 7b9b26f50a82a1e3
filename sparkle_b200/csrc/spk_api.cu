/*
 * spk_api.cu — the C ABI of include/spk_cuda.h: context lifetime, the device scene builder, camera, accumulation,
 * ray-batch entry points. Every entry point cites the reference interface it stands for in include/spk_cuda.h.
 * No CPU fallback: without a CUDA device spk_cuda_create fails (SPK_ERR_NO_DEVICE).
 */
#include "spk_context.h"

#include <cfloat>
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
template <typename Fn> void parallel_for(size_t n, Fn &&fn)
{
    unsigned workers = std::min<size_t>(std::min(8u, std::max(1u, std::thread::hardware_concurrency())), n);
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
    uint64_t leaf_prims = 0;
    while (!todo.empty())
    {
        auto [id, depth] = todo.back();
        todo.pop_back();
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

// nodes[1..] -> 64-byte pairs appended at pair slot `pair_base`; leaf words offset by `leaf_base`
void flatten_pairs(const SpkBvhDesc &d, uint32_t pair_base, uint32_t leaf_base, float *dst)
{
    auto rewrite = [&](uint32_t index) {
        uint32_t count = index & 15u, first = index >> 4;
        return count == 0 ? ((pair_base + (first - 1) / 2) << 4) : (((leaf_base + first) << 4) | count);
    };
    for (uint32_t i = 1; i < d.num_nodes; i++)
    {
        HostNode n = read_node(d, i);
        float *o = dst + size_t(i - 1) * 8;
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

template <typename T> int upload(SpkContext *ctx, const T *src, size_t count, T **out, bool keep = true)
{
    *out = nullptr;
    if (count == 0)
    {
        return SPK_OK;
    }
    void *p = nullptr;
    SPK_CUDA_TRY(cudaMalloc(&p, count * sizeof(T)));
    if (keep)
    {
        ctx->scene_allocs.push_back(p);
    }
    SPK_CUDA_TRY(cudaMemcpyAsync(p, src, count * sizeof(T), cudaMemcpyHostToDevice, ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream)); // host pointers are borrowed for this call only
    *out = static_cast<T *>(p);
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

int upload_texture(SpkContext *ctx, const SpkTextureDesc &t, DTexture &out)
{
    size_t tb = texel_bytes(t.format);
    if (!t.pixels || t.width == 0 || t.height == 0 || tb == 0)
    {
        return invalid("texture descriptor");
    }
    uint8_t *p = nullptr;
    int rc = upload(ctx, static_cast<const uint8_t *>(t.pixels), size_t(t.width) * t.height * tb, &p);
    if (rc != SPK_OK)
    {
        return rc;
    }
    out.pixels = p;
    out.width = t.width;
    out.height = t.height;
    out.format = t.format;
    out.pad = 0;
    return SPK_OK;
}

void free_scene(SpkContext *ctx)
{
    for (void *p : ctx->scene_allocs)
    {
        cudaFree(p);
    }
    ctx->scene_allocs.clear();
    ctx->has_scene = false;
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
    SPK_CUDA_TRY(cudaStreamCreateWithFlags(&ctx->stream, cudaStreamNonBlocking));
    SPK_CUDA_TRY(cudaEventCreate(&ctx->ev0));
    SPK_CUDA_TRY(cudaEventCreate(&ctx->ev1));
    SPK_CUDA_TRY(cudaMalloc(&ctx->d_accum, size_t(info->width) * info->height * sizeof(float4)));
    SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_accum, 0, size_t(info->width) * info->height * sizeof(float4), ctx->stream));
    SPK_CUDA_TRY(cudaMalloc(&ctx->d_counters, 8 * sizeof(unsigned long long)));
    SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_counters, 0, 8 * sizeof(unsigned long long), ctx->stream));
    SPK_CUDA_TRY(cudaStreamSynchronize(ctx->stream));
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
    cudaStreamSynchronize(ctx->stream);
    free_scene(ctx);
    cudaFree(ctx->d_accum);
    cudaFree(ctx->d_counters);
    cudaFree(ctx->d_rays);
    cudaFree(ctx->d_hits);
    cudaFree(ctx->d_resolve);
    cudaFree(ctx->wf_block);
    cudaFree(ctx->d_gbuf);
    cudaEventDestroy(ctx->ev0);
    cudaEventDestroy(ctx->ev1);
    if (ctx->own_stream)
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
    default:
        return invalid("spk_cuda_set_option: unknown option");
    }
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

    // ---- validation (before any allocation) ----
    if ((s->num_meshes && !s->meshes) || (s->num_primitives && !s->primitives) || (s->num_materials && !s->materials) ||
        (s->num_textures && !s->textures))
    {
        return invalid("spk_cuda_upload_scene: null table");
    }
    uint64_t total_pairs = 0, total_tris = 0;
    std::vector<int> mesh_status(s->num_meshes, 0); // 0 ok, 1 missing arrays, 2 index out of range, 3 malformed BLAS
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
        if (validate_bvh(md.blas, md.num_faces, "") != SPK_OK)
        {
            mesh_status[m] = 3;
        }
    });
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        if (mesh_status[m] != 0)
        {
            return invalid(mesh_status[m] == 1   ? "spk_cuda_upload_scene: mesh needs vertices, normals, tangents, indices"
                           : mesh_status[m] == 2 ? "spk_cuda_upload_scene: vertex index out of range"
                                                 : "spk_cuda_upload_scene: malformed BLAS");
        }
        total_pairs += (s->meshes[m].blas.num_nodes - 1) / 2;
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
    if (s->num_primitives)
    {
        if ((rc = validate_bvh(s->tlas, s->num_primitives, "spk_cuda_upload_scene: malformed TLAS")) != SPK_OK)
        {
            return rc;
        }
        total_pairs += (s->tlas.num_nodes - 1) / 2;
    }
    if (total_pairs >= (1ull << 28) || total_tris >= (1ull << 28))
    {
        return invalid("spk_cuda_upload_scene: scene exceeds 2^28 pairs or triangles");
    }

    // ---- flatten BLAS pairs + triangles (leaf order) ----
    // uninitialised on purpose: every word is written below, and zero-filling 80 MB first costs tens of ms of page faults
    // on one thread (the parallel writers below take the faults instead). One spare pair so an all-leaf scene still has
    // a valid (never dereferenced) pointer.
    const size_t pair_floats = size_t(total_pairs) * 16 + 16, tri_floats = size_t(total_tris) * 12;
    std::unique_ptr<float[]> pairs_buf(new (std::nothrow) float[pair_floats]);
    std::unique_ptr<float[]> tris_buf(new (std::nothrow) float[tri_floats ? tri_floats : 1]);
    if (!pairs_buf || !tris_buf)
    {
        return SPK_ERR_OUT_OF_MEMORY;
    }
    memset(pairs_buf.get() + size_t(total_pairs) * 16, 0, 64);
    struct
    {
        float *p;
        float *data() const { return p; }
    } pairs{pairs_buf.get()}, tris{tris_buf.get()};
    std::vector<DMesh> dmeshes(s->num_meshes);
    ctx->mesh_root_word.assign(s->num_meshes, 0);
    uint32_t pair_base = 0, tri_base = 0;
    // work items: one flatten_pairs per mesh, and the triangle records in chunks of 64 Ki faces
    struct Item
    {
        uint32_t mesh, pair_base, tri_base, begin, end; // begin == end: the pair item
    };
    std::vector<Item> items;
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        const SpkMeshDesc &md = s->meshes[m];
        items.push_back(Item{m, pair_base, tri_base, 0, 0});
        for (uint32_t b = 0; b < md.num_faces; b += 65536u)
        {
            items.push_back(Item{m, pair_base, tri_base, b, std::min(md.num_faces, b + 65536u)});
        }
        pair_base += (md.blas.num_nodes - 1) / 2;
        tri_base += md.num_faces;
    }
    parallel_for(items.size(), [&](size_t w) {
        const Item &it = items[w];
        const SpkMeshDesc &md = s->meshes[it.mesh];
        if (it.begin == it.end)
        {
            flatten_pairs(md.blas, it.pair_base, it.tri_base, pairs.data() + size_t(it.pair_base) * 16);
            ctx->mesh_root_word[it.mesh] = root_word(md.blas, it.pair_base, it.tri_base);
            return;
        }
        for (uint32_t k = it.begin; k < it.end; k++)
        {
            uint32_t face = read_prim_id(md.blas, k);
            const float *v0 = md.vertices + size_t(md.indices[size_t(face) * 3 + 0]) * 3;
            const float *v1 = md.vertices + size_t(md.indices[size_t(face) * 3 + 1]) * 3;
            const float *v2 = md.vertices + size_t(md.indices[size_t(face) * 3 + 2]) * 3;
            float *o = tris.data() + size_t(it.tri_base + k) * 12;
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
    });
    pair_base = 0;
    tri_base = 0;
    for (uint32_t m = 0; m < s->num_meshes; m++)
    {
        const SpkMeshDesc &md = s->meshes[m];
        float *dn = nullptr, *dt = nullptr, *du = nullptr;
        uint32_t *di = nullptr;
        if ((rc = upload(ctx, md.normals, size_t(md.num_vertices) * 3, &dn)) != SPK_OK ||
            (rc = upload(ctx, md.tangents, size_t(md.num_vertices) * 4, &dt)) != SPK_OK ||
            (rc = upload(ctx, md.indices, size_t(md.num_faces) * 3, &di)) != SPK_OK ||
            (md.uvs && (rc = upload(ctx, md.uvs, size_t(md.num_vertices) * 2, &du)) != SPK_OK))
        {
            free_scene(ctx);
            return rc;
        }
        dmeshes[m].normals = dn;
        dmeshes[m].tangents = dt;
        dmeshes[m].uvs = du;
        dmeshes[m].indices = di;
        pair_base += (md.blas.num_nodes - 1) / 2;
        tri_base += md.num_faces;
    }
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
    std::vector<DInstance> instances(s->num_primitives);
    if (s->num_primitives)
    {
        flatten_pairs(s->tlas, pair_base, 0, pairs.data() + size_t(pair_base) * 16);
        ctx->scene.tlas_root = root_word(s->tlas, pair_base, 0);
        for (uint32_t k = 0; k < s->num_primitives; k++)
        {
            uint32_t prim = read_prim_id(s->tlas, k);
            fill_instance(instances[k], s->primitives[prim], prim, ctx->mesh_root_word, ctx->material_class);
        }
    }
    ctx->pair_capacity = uint32_t(total_pairs);
    ctx->instance_capacity = s->num_primitives;

    float4 *d_pairs = nullptr, *d_tris = nullptr;
    DInstance *d_inst = nullptr;
    DMesh *d_meshes = nullptr;
    if ((rc = upload(ctx, reinterpret_cast<const float4 *>(pairs.data()), pair_floats / 4, &d_pairs)) != SPK_OK ||
        (rc = upload(ctx, reinterpret_cast<const float4 *>(tris.data()), tri_floats / 4, &d_tris)) != SPK_OK ||
        (rc = upload(ctx, instances.data(), instances.size(), &d_inst)) != SPK_OK ||
        (rc = upload(ctx, dmeshes.data(), dmeshes.size(), &d_meshes)) != SPK_OK)
    {
        free_scene(ctx);
        return rc;
    }
    ctx->d_pairs = d_pairs;
    ctx->d_instances = d_inst;

    // ---- materials, textures, sky ----
    std::vector<DTexture> dtex(s->num_textures);
    for (uint32_t t = 0; t < s->num_textures; t++)
    {
        if ((rc = upload_texture(ctx, s->textures[t], dtex[t])) != SPK_OK)
        {
            free_scene(ctx);
            return rc;
        }
    }
    static_assert(sizeof(SpkMaterialDesc) == sizeof(DMaterial), "material ABI == device record");
    DMaterial *d_mats = nullptr;
    DTexture *d_tex = nullptr;
    if ((rc = upload(ctx, reinterpret_cast<const DMaterial *>(s->materials), s->num_materials, &d_mats)) != SPK_OK ||
        (rc = upload(ctx, dtex.data(), dtex.size(), &d_tex)) != SPK_OK)
    {
        free_scene(ctx);
        return rc;
    }
    DSky sky{};
    sky.has_sky = s->sky.has_sky;
    sky.has_map = s->sky.has_sky ? s->sky.has_map : 0;
    memcpy(sky.color, s->sky.color, 12);
    if (sky.has_map)
    {
        for (int f = 0; f < 6; f++)
        {
            if ((rc = upload_texture(ctx, s->sky.faces[f], sky.faces[f])) != SPK_OK)
            {
                free_scene(ctx);
                return rc;
            }
        }
    }

    ctx->scene.pairs = d_pairs;
    ctx->scene.tris = d_tris;
    ctx->scene.instances = d_inst;
    ctx->scene.meshes = d_meshes;
    ctx->scene.materials = d_mats;
    ctx->scene.textures = d_tex;
    ctx->scene.num_instances = s->num_primitives;
    ctx->scene.cull_extent = s->world_extent > 0.f ? s->world_extent : 1.f;
    ctx->scene.sky = sky;
    ctx->has_scene = true;
    ctx->tracer_choice = -1; // measured again for the new scene (spk_cuda_choose_tracer)
    return SPK_OK;
}

int spk_cuda_update_instances(SpkContext *ctx, const SpkPrimitiveDesc *primitives, uint32_t num_primitives,
                              const SpkBvhDesc *tlas)
{
    if (!ctx || !primitives || !tlas)
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
    // the t-culling slack is sized from the world extent (spk_traverse.cuh): primitives may have moved outside the
    // bounds of the upload, so take it again from the rebuilt TLAS root (= the world AABB)
    const HostNode root = read_node(*tlas, 0);
    float extent = 0.f;
    for (float b : root.bounds)
    {
        extent = std::max(extent, std::fabs(b));
    }
    if (extent > 0.f && std::isfinite(extent))
    {
        ctx->scene.cull_extent = extent;
    }
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
    SPK_CUDA_TRY(cudaSetDevice(ctx->device));
    SPK_CUDA_TRY(cudaMemsetAsync(ctx->d_accum, 0, size_t(ctx->width) * ctx->height * sizeof(float4), ctx->stream));
    ctx->accum_samples = 0;
    return SPK_OK;
}

int spk_cuda_accum_device_ptr(SpkContext *ctx, void **device_ptr, size_t *bytes)
{
    if (!ctx || !device_ptr || !bytes)
    {
        return invalid("spk_cuda_accum_device_ptr: null argument");
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
