/* spk_context.h — internal state behind the opaque SpkContext of include/spk_cuda.h (not part of the ABI). */
#pragma once
#include "../../include/spk_cuda.h"
#include "spk_device.cuh"

#include <algorithm>
#include <cstdlib>
#include <cuda_runtime.h>
#include <string>
#include <vector>

struct SpkContext
{
    int device = 0;
    cudaStream_t stream = nullptr;
    bool own_stream = true;
    uint32_t width = 0, height = 0;

    bool has_scene = false;
    spk::DScene scene{};
    // every table of the current scene lives in one device arena, filled from one pinned staging buffer; both are kept
    // across uploads and only grow (spk_api.cu)
    void *d_arena = nullptr;
    size_t arena_capacity = 0;
    void *h_stage = nullptr;
    size_t stage_capacity = 0;
    // kept for spk_cuda_update_instances
    std::vector<uint32_t> mesh_root_word;
    std::vector<uint32_t> material_class; // per material: 2 dielectric + 4 textured + 8 emissive (DInstance::shade_class)
    uint32_t tlas_pair_base = 0;
    uint32_t num_blas_pairs = 0;
    uint32_t pair_capacity = 0;
    uint32_t instance_capacity = 0;
    float4 *d_pairs = nullptr;
    spk::DInstance *d_instances = nullptr;

    bool has_camera = false;
    spk::DCamera camera{};

    float4 *d_accum = nullptr; // W*H float4
    float4 *d_gbuf = nullptr;  // 2 x W*H float4 (colour + valid, first-hit normal), allocated on first denoised render
    uint32_t accum_samples = 0;
    bool accum_clear_pending = false; // reset requested; done by the next megakernel launch for free, else by a memset
    uint32_t accum_mode = 0; // SpkAccumMode of the samples in d_accum

    unsigned long long *d_counters = nullptr; // 8 x u64
    cudaEvent_t ev0 = nullptr, ev1 = nullptr;
    SpkStats stats{};

    // scratch for the host-buffer ray-batch entry point
    void *d_rays = nullptr;
    void *d_hits = nullptr;
    size_t ray_capacity = 0;
    // pinned host buffers: copy-in / copy-out streams and events of the three-stage pipeline (spk_cuda_trace_closest)
    cudaStream_t copy_in_stream = nullptr, copy_out_stream = nullptr;
    cudaEvent_t pipe_ev[16] = {};

    void *d_resolve = nullptr;
    size_t resolve_capacity = 0;
    void *h_resolve = nullptr; // pinned bounce buffer for pageable destinations (spk_cuda_resolve)
    size_t h_resolve_capacity = 0;
    cudaEvent_t resolve_ev[4] = {nullptr, nullptr, nullptr, nullptr};

    // device BVH builder (spk_build.cu): scratch kept across builds; build_ids = sorted primitive ids of the last build
    void *d_build_scratch = nullptr;
    size_t build_scratch_capacity = 0;
    uint32_t *build_ids = nullptr;
    float *d_tlas_boxes = nullptr; // world boxes of the primitives, in the scene arena
    // BLAS leaves of device-built trees (the TLAS always gets one primitive per leaf). An LBVH's leaves are spatially
    // loose: with 4 triangles per leaf soup-1M costs 13.7 triangle tests per ray against 2.1 with one, 1243 vs 1648 Mrays/s
    uint32_t opt_build_max_leaf = 1;
    // where every tree of the current scene lives (spk_cuda_download_bvh, spk_cuda_update_instances, build reports)
    struct TreeInfo
    {
        uint32_t pair_base = 0, num_pairs = 0, leaf_base = 0, num_prims = 0, root_word = 0, depth = 0;
        float root_bounds[6] = {0, 0, 0, 0, 0, 0}; // min.xyz, max.xyz
        double sah_cost = 0.0;
        double build_ms = 0.0;
        bool device_built = false;
    };
    std::vector<TreeInfo> mesh_trees;
    TreeInfo tlas_tree;
    // sorted megakernel: strips of tiles launched in the order of their cost in the previous launch of the view
    uint32_t *d_strip = nullptr;
    uint32_t strip_capacity = 0, strip_order_count = 0;
    size_t strip_tmp_bytes = 0;
    int strip_order_form = 0;
    bool strip_order_valid = false;
    bool opt_cost_order = true;
    void *d_group_block = nullptr; // multi-GPU: flags + the root's gathered output, visible to peers (spk_group.cu)

    // wavefront tracer: path-slot pool (spk_wavefront.cu)
    void *wf_block = nullptr;
    uint32_t wf_slots = 0;
    size_t wf_bytes = 0;
    int mega_variant = 1084; // megakernel form (SPK_OPT_MEGAKERNEL_FORM code) spk_cuda_choose_tracer measured fastest
    // explicit options (spk_cuda_set_option); nothing is read from the environment
    int opt_mega_form = 0;           // 0: use mega_variant
    uint32_t opt_wf_slots = 1u << 26; // 64 Mi slots = 7.2 GB (capped by the work of the call): fewer launch tails; city-10M at 4K x 32 spp: 4 / 8 / 16 / 32 / 64 Mi = 398 / 414 / 425 / 434 / 442 Mpaths/s
    uint32_t opt_wf_chunk = 8;
    uint32_t opt_wf_sort = 0; // wavefront ray sorting: bits per axis of the origin cell in the sort key (0 = off)
    int opt_vote = 6;
    int opt_trace_top = 0; // pairs of the hottest tree's top staged in shared memory by the ray-batch kernel (0 = off)
    int opt_carveout = -2;
    bool opt_verbose = false;
    int tracer_choice = -1; // -1 not measured for this scene, 0 megakernel, 1 wavefront (spk_cuda_choose_tracer)
};

namespace spk
{
void set_error(const std::string &msg);
int cuda_fail(cudaError_t e, const char *what);
#define SPK_CUDA_TRY(expr)                                                                                             \
    do                                                                                                                 \
    {                                                                                                                  \
        cudaError_t e__ = (expr);                                                                                      \
        if (e__ != cudaSuccess)                                                                                        \
            return spk::cuda_fail(e__, #expr);                                                                         \
    } while (0)

// Shared memory and L1 split one 256 KB array per SM. Asks for the smallest carve-out that still holds `ctas` resident
// CTAs of `bytes_per_cta` (+1 KB the driver reserves per CTA), so that the rest serves the BVH as L1: the default
// heuristic leaves the traversal kernels ~32 KB less L1 than they can have (measured: DESIGN.md 4.2).
template <class Kernel>
cudaError_t prefer_small_carveout(Kernel kernel, size_t bytes_per_cta, int ctas, int override_pct = -2)
{
    const size_t pool = 228 * 1024;
    int pct = int((size_t(ctas) * (bytes_per_cta + 1024) * 100 + pool - 1) / pool);
    if (override_pct != -2)
        pct = override_pct; // SPK_OPT_SMEM_CARVEOUT
    if (pct < 0)
        return cudaSuccess; // -1: the driver's default
    return cudaFuncSetAttribute(kernel, cudaFuncAttributePreferredSharedMemoryCarveout, pct > 100 ? 100 : pct);
}

// clears the accumulator now if spk_cuda_reset_accumulation left it pending (consumers other than the megakernels)
int ensure_accum_cleared(SpkContext *ctx);
struct DeviceBuildResult
{
    uint32_t num_pairs, root_word, depth;
    float root_bounds[6]; // min.xyz, max.xyz
    double sah_cost;
};
// spk_build.cu: LBVH over triangles (d_vertices / d_indices; leaf-ordered records to d_tris) or over boxes (d_boxes_in)
int device_build_lbvh(SpkContext *ctx, const float *d_vertices, const uint32_t *d_indices, const float *d_boxes_in, uint32_t n,
                      uint32_t max_leaf, float4 *d_pairs, uint32_t pair_base, uint32_t leaf_base, float4 *d_tris,
                      DeviceBuildResult *out);
// kernels (spk_trace.cu / spk_render.cu)
int launch_trace_closest(SpkContext *ctx, const void *d_rays, uint64_t n, void *d_hits, uint32_t flags);
int launch_render(SpkContext *ctx, const SpkRenderParams *params);
int launch_render_wavefront(SpkContext *ctx, const SpkRenderParams *params);
int launch_denoise_accumulate(SpkContext *ctx, uint32_t spatial, uint32_t accum_mode, uint32_t cumulated);
int copy_to_host(SpkContext *ctx, void *host_out, const void *d_src, size_t bytes);
} // namespace spk
