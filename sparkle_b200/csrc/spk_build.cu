/*
 * spk_build.cu — device-side BVH builder: TLAS::Build / BLAS::Build on the GPU.
 *
 * The reference builds both levels on the host with bvh::v2::DefaultBuilder (SceneRenderProxy.cpp:31-69,
 * MeshRenderProxy.cpp:31-65); every scene of round 1 waited on a host SAH build. Here a mesh whose SpkBvhDesc has
 * nodes == NULL is built on the device, straight into the 64-byte pair layout the traversal kernels read:
 *
 *   boxes + centroids  -> 30-bit Morton codes on the centroid bounds -> radix sort (cub::DeviceRadixSort, the one library
 *   call: a plain key-value sort) -> Karras' binary radix tree over the sorted codes (ties broken by position) -> bottom-up
 *   AABB fit -> subtrees of <= max_leaf primitives collapse into leaves (sorted order makes every subtree a contiguous run)
 *   -> surviving inner nodes numbered in depth-first preorder (a prefix sum + one walk to the root per node: parents precede
 *   children, a left child's pair is the next 64 bytes) -> sibling pairs and leaf-ordered triangle records written in place.
 *
 * The same kernels build the TLAS from the instances' world boxes (max_leaf 1). spk_cuda_download_bvh hands any tree on the
 * device back in the bvh::v2 node layout + prim_ids, so the oracle traverses the SAME topology (tests/test_build_gpu.py:
 * closest hits bit-exact against the reference on the device-built tree, and equal to brute force).
 * Leaves hold one triangle by default (SPK_OPT_BUILD_MAX_LEAF): an LBVH's leaves are spatially loose, larger ones cost more
 * triangle tests than they save node visits. An LBVH is a fast build, not a SAH-optimal one: the host builder (host/bvh_builder.cpp, binned SAH) stays the default for
 * trees handed in by the caller; spk_cuda_get_build_info reports the SAH cost of either for comparison.
 * The reference's traversal stack holds 32 entries per level: a device-built tree deeper than 32 is refused.
 */
#include "spk_context.h"

#include <cfloat>
#include <cub/device/device_radix_sort.cuh>
#include <cub/device/device_scan.cuh>
#include <cstring>

namespace spk
{
namespace
{
constexpr uint32_t kLeafBit = 0x80000000u;

struct BuildBuffers // all sized by n, carved from one scratch allocation
{
    float *boxes;      // 6 n: min.xyz, max.xyz
    uint32_t *codes, *codes_sorted;
    uint32_t *ids, *ids_sorted;
    uint32_t *left, *right, *parent; // per internal node (n - 1); child: index | kLeafBit for a sorted-primitive leaf
    uint32_t *first, *last;          // sorted-position range of every internal node
    float *node_boxes;               // 6 (n - 1)
    uint32_t *visits;                // bottom-up fit counters
    uint32_t *inner, *slot;          // survives as an inner node; its pair slot
    uint32_t *leaf_parent;           // parent of every sorted-primitive leaf
    float *bounds;                   // 6: centroid bounds (ordered-int encoded while reducing)
    uint32_t *stats;                 // [0] max depth, [1] leaves, [2..3] SAH cost accumulators (float bits)
};

__device__ __forceinline__ int ordered(float f)
{
    int i = __float_as_int(f);
    return i >= 0 ? i : i ^ 0x7FFFFFFF;
}
__device__ __forceinline__ float unordered(int i)
{
    return __int_as_float(i >= 0 ? i : i ^ 0x7FFFFFFF);
}

__global__ void init_bounds_kernel(int *bounds)
{
    if (threadIdx.x < 3)
    {
        bounds[threadIdx.x] = ordered(FLT_MAX);
        bounds[3 + threadIdx.x] = ordered(-FLT_MAX);
    }
}

// Mesh::GetTriangleMinMax (io/Mesh.h:55-71): min / max of the three vertices
__global__ void triangle_boxes_kernel(const float *__restrict__ vertices, const uint32_t *__restrict__ indices, uint32_t n,
                                      float *__restrict__ boxes, int *bounds)
{
    const uint32_t f = blockIdx.x * blockDim.x + threadIdx.x;
    float c[3] = {0.f, 0.f, 0.f};
    const bool live = f < n;
    if (live)
    {
        const float *v0 = vertices + size_t(indices[size_t(f) * 3]) * 3;
        const float *v1 = vertices + size_t(indices[size_t(f) * 3 + 1]) * 3;
        const float *v2 = vertices + size_t(indices[size_t(f) * 3 + 2]) * 3;
        for (int a = 0; a < 3; a++)
        {
            const float lo = fminf(v0[a], fminf(v1[a], v2[a])), hi = fmaxf(v0[a], fmaxf(v1[a], v2[a]));
            boxes[size_t(f) * 6 + a] = lo;
            boxes[size_t(f) * 6 + 3 + a] = hi;
            c[a] = (lo + hi) / 2.f; // MeshRenderProxy.cpp:50
        }
    }
    for (int a = 0; a < 3; a++)
    {
        float lo = live ? c[a] : FLT_MAX, hi = live ? c[a] : -FLT_MAX;
        for (int d = 16; d; d >>= 1)
        {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
        }
        if ((threadIdx.x & 31) == 0 && lo <= hi)
        {
            atomicMin(bounds + a, ordered(lo));
            atomicMax(bounds + 3 + a, ordered(hi));
        }
    }
}

// boxes given (TLAS): only the centroid bounds
__global__ void box_bounds_kernel(const float *__restrict__ boxes, uint32_t n, int *bounds)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    const bool live = k < n;
    for (int a = 0; a < 3; a++)
    {
        const float c = live ? (boxes[size_t(k) * 6 + a] + boxes[size_t(k) * 6 + 3 + a]) / 2.f : 0.f;
        float lo = live ? c : FLT_MAX, hi = live ? c : -FLT_MAX;
        for (int d = 16; d; d >>= 1)
        {
            lo = fminf(lo, __shfl_xor_sync(0xffffffffu, lo, d));
            hi = fmaxf(hi, __shfl_xor_sync(0xffffffffu, hi, d));
        }
        if ((threadIdx.x & 31) == 0 && lo <= hi)
        {
            atomicMin(bounds + a, ordered(lo));
            atomicMax(bounds + 3 + a, ordered(hi));
        }
    }
}

__device__ __forceinline__ uint32_t expand_bits(uint32_t v)
{
    v = (v * 0x00010001u) & 0xFF0000FFu;
    v = (v * 0x00000101u) & 0x0F00F00Fu;
    v = (v * 0x00000011u) & 0xC30C30C3u;
    v = (v * 0x00000005u) & 0x49249249u;
    return v;
}

__global__ void morton_kernel(const float *__restrict__ boxes, uint32_t n, const int *__restrict__ bounds,
                              uint32_t *__restrict__ codes, uint32_t *__restrict__ ids)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n)
    {
        return;
    }
    uint32_t q[3];
    for (int a = 0; a < 3; a++)
    {
        const float lo = unordered(bounds[a]), hi = unordered(bounds[3 + a]);
        const float c = (boxes[size_t(k) * 6 + a] + boxes[size_t(k) * 6 + 3 + a]) / 2.f;
        const float ext = hi - lo;
        float t = ext > 0.f ? (c - lo) / ext : 0.f;
        t = fminf(fmaxf(t * 1024.f, 0.f), 1023.f);
        q[a] = (uint32_t)t;
    }
    codes[k] = (expand_bits(q[0]) << 2) | (expand_bits(q[1]) << 1) | expand_bits(q[2]);
    ids[k] = k;
}

// common-prefix length of sorted positions i and j; equal codes fall back to the positions themselves
__device__ __forceinline__ int delta(const uint32_t *__restrict__ codes, int n, int i, int j)
{
    if (j < 0 || j >= n)
    {
        return -1;
    }
    const uint32_t a = codes[i], b = codes[j];
    return a != b ? __clz(a ^ b) : 32 + __clz((uint32_t)i ^ (uint32_t)j);
}

// Karras 2012: one thread per internal node of the binary radix tree
__global__ void radix_tree_kernel(const uint32_t *__restrict__ codes, int n, uint32_t *__restrict__ left,
                                  uint32_t *__restrict__ right, uint32_t *__restrict__ parent,
                                  uint32_t *__restrict__ leaf_parent, uint32_t *__restrict__ first_out,
                                  uint32_t *__restrict__ last_out)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1)
    {
        return;
    }
    const int d = delta(codes, n, i, i + 1) > delta(codes, n, i, i - 1) ? 1 : -1;
    const int dmin = delta(codes, n, i, i - d);
    int lmax = 2;
    while (delta(codes, n, i, i + lmax * d) > dmin)
    {
        lmax *= 2;
    }
    int l = 0;
    for (int t = lmax / 2; t >= 1; t /= 2)
    {
        if (delta(codes, n, i, i + (l + t) * d) > dmin)
        {
            l += t;
        }
    }
    const int j = i + l * d;
    const int dnode = delta(codes, n, i, j);
    int s = 0;
    for (int t = (l + 1) / 2;; t = (t + 1) / 2)
    {
        if (delta(codes, n, i, i + (s + t) * d) > dnode)
        {
            s += t;
        }
        if (t == 1)
        {
            break;
        }
    }
    const int gamma = i + s * d + min(d, 0);
    const int lo = min(i, j), hi = max(i, j);
    const uint32_t lc = lo == gamma ? (uint32_t)gamma | kLeafBit : (uint32_t)gamma;
    const uint32_t rc = hi == gamma + 1 ? (uint32_t)(gamma + 1) | kLeafBit : (uint32_t)(gamma + 1);
    left[i] = lc;
    right[i] = rc;
    first_out[i] = (uint32_t)lo;
    last_out[i] = (uint32_t)hi;
    if (lc & kLeafBit)
    {
        leaf_parent[gamma] = (uint32_t)i;
    }
    else
    {
        parent[gamma] = (uint32_t)i;
    }
    if (rc & kLeafBit)
    {
        leaf_parent[gamma + 1] = (uint32_t)i;
    }
    else
    {
        parent[gamma + 1] = (uint32_t)i;
    }
    if (i == 0)
    {
        parent[0] = 0xFFFFFFFFu;
    }
}

// bottom-up: the second thread to reach a node owns it (both children are complete)
__global__ void fit_bounds_kernel(const float *__restrict__ boxes, const uint32_t *__restrict__ ids, int n,
                                  const uint32_t *__restrict__ left, const uint32_t *__restrict__ right,
                                  const uint32_t *__restrict__ parent, const uint32_t *__restrict__ leaf_parent,
                                  float *__restrict__ node_boxes, uint32_t *__restrict__ visits)
{
    const int k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n)
    {
        return;
    }
    uint32_t node = leaf_parent[k];
    while (node != 0xFFFFFFFFu)
    {
        if (atomicAdd(&visits[node], 1u) == 0u)
        {
            return; // first arrival: the sibling subtree is not finished yet
        }
        __threadfence();
        float b[6];
        for (int side = 0; side < 2; side++)
        {
            const uint32_t c = side ? right[node] : left[node];
            const float *src = (c & kLeafBit) ? boxes + size_t(ids[c & ~kLeafBit]) * 6 : node_boxes + size_t(c) * 6;
            for (int a = 0; a < 3; a++)
            {
                // plain loads would be fine for leaf boxes; inner boxes were written by another thread just now
                const float lo = __ldcg(src + a), hi = __ldcg(src + 3 + a);
                b[a] = side ? fminf(b[a], lo) : lo;
                b[3 + a] = side ? fmaxf(b[3 + a], hi) : hi;
            }
        }
        for (int a = 0; a < 6; a++)
        {
            __stcg(node_boxes + size_t(node) * 6 + a, b[a]);
        }
        __threadfence();
        node = parent[node];
    }
}

// which radix-tree nodes survive as inner nodes (more than max_leaf primitives below them)
__global__ void mark_inner_kernel(const uint32_t *__restrict__ first, const uint32_t *__restrict__ last, int n,
                                  uint32_t max_leaf, uint32_t *__restrict__ inner)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n - 1)
    {
        inner[i] = (last[i] - first[i] + 1u) > max_leaf ? 1u : 0u;
    }
}

// Pair slots in DEPTH-FIRST PREORDER: a parent precedes its children (the bvh::v2 layout of the host builder, and what
// validate_bvh checks) and the left child's pair is the next 64 bytes, i.e. often the other half of the 128-byte line
// the parent's fetch just brought in. The internal nodes of a subtree are a contiguous run of indices, so with S = exclusive
// prefix sum of the `inner` flags the number of surviving inner nodes in a subtree is a difference of two entries of S, and
// preorder(node) = sum over the path to the root of (1 + inner nodes of the left sibling subtree when
// the path goes right; a left child [f, g] has index g and owns the internal indices f+1 .. g). A level-by-level numbering
// (sort by depth) was measured too: the same speed on incoherent rays (soup-1M 1238 vs 1236 Mrays/s), so the order that
// matches the host builder's was kept.
__global__ void preorder_slots_kernel(const uint32_t *__restrict__ left, const uint32_t *__restrict__ parent,
                                      const uint32_t *__restrict__ first, const uint32_t *__restrict__ last,
                                      const uint32_t *__restrict__ inner, const uint32_t *__restrict__ prefix, int n,
                                      uint32_t *__restrict__ slot, uint32_t *__restrict__ stats)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    const bool is_inner = i < n - 1 && inner[i];
    uint32_t depth = 0;
    if (is_inner)
    {
        uint32_t pre = 0, c = (uint32_t)i;
        for (uint32_t p = parent[c]; p != 0xFFFFFFFFu; c = p, p = parent[p])
        {
            pre += 1u;
            const uint32_t l = left[p];
            if (l != c && !(l & kLeafBit))
            {
                // surviving inner nodes of the left sibling's subtree: a LEFT child sits at the last position of its range
                // [f, g], and the internal nodes below it are f+1 .. g
                pre += prefix[last[l] + 1u] - prefix[first[l] + 1u];
            }
            depth++;
        }
        slot[i] = pre;
    }
    uint32_t deepest = is_inner ? depth + 1u : 0u; // its children sit one level below
    for (int d = 16; d; d >>= 1)
    {
        deepest = max(deepest, __shfl_xor_sync(0xffffffffu, deepest, d));
    }
    if ((threadIdx.x & 31) == 0 && deepest)
    {
        atomicMax(&stats[0], deepest);
    }
}

__device__ __forceinline__ float half_area(const float *b)
{
    const float dx = b[3] - b[0], dy = b[4] - b[1], dz = b[5] - b[2];
    return dx * dy + dy * dz + dz * dx;
}

// one thread per surviving inner node: its two children as one 64-byte pair record, index words absolute
__global__ void emit_pairs_kernel(const float *__restrict__ boxes, const uint32_t *__restrict__ ids, int n,
                                  const uint32_t *__restrict__ left, const uint32_t *__restrict__ right,
                                  const uint32_t *__restrict__ parent, const uint32_t *__restrict__ first,
                                  const uint32_t *__restrict__ last, const float *__restrict__ node_boxes,
                                  const uint32_t *__restrict__ inner, const uint32_t *__restrict__ slot,
                                  uint32_t pair_base, uint32_t leaf_base, float4 *__restrict__ pairs,
                                  uint32_t *__restrict__ stats)
{
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= n - 1 || !inner[i])
    {
        return;
    }
    float4 rec[4];
    uint32_t leaves = 0;
    float cost = 0.f;
    for (int side = 0; side < 2; side++)
    {
        const uint32_t c = side ? right[i] : left[i];
        const float *b;
        uint32_t word;
        if (c & kLeafBit)
        {
            const uint32_t pos = c & ~kLeafBit;
            b = boxes + size_t(ids[pos]) * 6;
            word = ((leaf_base + pos) << 4) | 1u;
            leaves++;
            cost += half_area(b);
        }
        else
        {
            b = node_boxes + size_t(c) * 6;
            if (inner[c])
            {
                word = (pair_base + slot[c]) << 4;
            }
            else
            {
                const uint32_t count = last[c] - first[c] + 1u;
                word = ((leaf_base + first[c]) << 4) | count;
                leaves++;
                cost += half_area(b) * (float)count;
            }
        }
        // bvh::v2::Node bounds order: min.x, max.x, min.y, max.y, min.z, max.z
        rec[2 * side] = make_float4(b[0], b[3], b[1], b[4]);
        rec[2 * side + 1] = make_float4(b[2], b[5], __uint_as_float(word), 0.f);
    }
    float4 *dst = pairs + size_t(pair_base + slot[i]) * 4;
    dst[0] = rec[0];
    dst[1] = rec[1];
    dst[2] = rec[2];
    dst[3] = rec[3];
    (void)parent;
    (void)leaves;
    // SAH cost with unit costs: sum of inner areas + sum of leaf area * count, over the root area (done on the host)
    atomicAdd(reinterpret_cast<float *>(&stats[2]), half_area(node_boxes + size_t(i) * 6) + cost);
}

// leaf-ordered triangle records: Triangle::Set (MeshRenderProxy.cpp:80-86), the same float subtractions as the host path
__global__ void emit_triangles_kernel(const float *__restrict__ vertices, const uint32_t *__restrict__ indices,
                                      const uint32_t *__restrict__ ids, uint32_t n, float4 *__restrict__ tris)
{
    const uint32_t k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= n)
    {
        return;
    }
    const uint32_t face = ids[k];
    const float *v0 = vertices + size_t(indices[size_t(face) * 3]) * 3;
    const float *v1 = vertices + size_t(indices[size_t(face) * 3 + 1]) * 3;
    const float *v2 = vertices + size_t(indices[size_t(face) * 3 + 2]) * 3;
    float4 *o = tris + size_t(k) * 3;
    o[0] = make_float4(v0[0], v0[1], v0[2], __uint_as_float(face));
    o[1] = make_float4(v1[0] - v0[0], v1[1] - v0[1], v1[2] - v0[2], 0.f);
    o[2] = make_float4(v2[0] - v0[0], v2[1] - v0[1], v2[2] - v0[2], 0.f);
}

size_t align_up(size_t v)
{
    return (v + 255u) & ~size_t(255u);
}
} // namespace

// Builds a tree over n primitives on the context's stream. Triangles: d_vertices / d_indices given, boxes computed here and
// the leaf-ordered triangle records written to d_tris (n records). Boxes: d_boxes_in given (6 floats each), no triangles.
// Pairs go to d_pairs[pair_base ..]; leaf index words are offset by leaf_base. The sorted primitive ids stay in
// ctx->build_ids (device) for the caller. Blocking (the result sizes come back to the host).
int device_build_lbvh(SpkContext *ctx, const float *d_vertices, const uint32_t *d_indices, const float *d_boxes_in, uint32_t n,
                      uint32_t max_leaf, float4 *d_pairs, uint32_t pair_base, uint32_t leaf_base, float4 *d_tris,
                      DeviceBuildResult *out)
{
    memset(out, 0, sizeof(*out));
    if (n == 0 || max_leaf == 0 || max_leaf > 15)
    {
        set_error("device BVH build: empty input or max_leaf outside [1, 15]");
        return SPK_ERR_INVALID_ARGUMENT;
    }
    cudaStream_t s = ctx->stream;
    // ---- scratch: one allocation, kept and grown across builds ----
    const size_t nn = n;
    size_t sort_bytes = 0, scan_bytes = 0;
    cub::DeviceRadixSort::SortPairs(nullptr, sort_bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr,
                                    (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, 0, 30, s);
    cub::DeviceScan::ExclusiveSum(nullptr, scan_bytes, (const uint32_t *)nullptr, (uint32_t *)nullptr, (int)n, s);
    size_t need = 0;
    auto take = [&](size_t bytes) {
        const size_t at = need;
        need += align_up(bytes);
        return at;
    };
    const size_t o_boxes = take(nn * 24), o_codes = take(nn * 4), o_codes2 = take(nn * 4), o_ids = take(nn * 4),
                 o_ids2 = take(nn * 4), o_left = take(nn * 4), o_right = take(nn * 4), o_parent = take(nn * 4),
                 o_first = take(nn * 4), o_last = take(nn * 4), o_nboxes = take(nn * 24), o_visits = take(nn * 4),
                 o_inner = take(nn * 4), o_slot = take(nn * 4), o_lparent = take(nn * 4), o_bounds = take(64),
                 o_stats = take(64), o_cub = take(std::max(sort_bytes, scan_bytes));
    if (ctx->build_scratch_capacity < need)
    {
        cudaFree(ctx->d_build_scratch);
        ctx->d_build_scratch = nullptr;
        ctx->build_scratch_capacity = 0;
        SPK_CUDA_TRY(cudaMalloc(&ctx->d_build_scratch, need));
        ctx->build_scratch_capacity = need;
    }
    uint8_t *base = static_cast<uint8_t *>(ctx->d_build_scratch);
    BuildBuffers b;
    b.boxes = reinterpret_cast<float *>(base + o_boxes);
    b.codes = reinterpret_cast<uint32_t *>(base + o_codes);
    b.codes_sorted = reinterpret_cast<uint32_t *>(base + o_codes2);
    b.ids = reinterpret_cast<uint32_t *>(base + o_ids);
    b.ids_sorted = reinterpret_cast<uint32_t *>(base + o_ids2);
    b.left = reinterpret_cast<uint32_t *>(base + o_left);
    b.right = reinterpret_cast<uint32_t *>(base + o_right);
    b.parent = reinterpret_cast<uint32_t *>(base + o_parent);
    b.first = reinterpret_cast<uint32_t *>(base + o_first);
    b.last = reinterpret_cast<uint32_t *>(base + o_last);
    b.node_boxes = reinterpret_cast<float *>(base + o_nboxes);
    b.visits = reinterpret_cast<uint32_t *>(base + o_visits);
    b.inner = reinterpret_cast<uint32_t *>(base + o_inner);
    b.slot = reinterpret_cast<uint32_t *>(base + o_slot);
    b.leaf_parent = reinterpret_cast<uint32_t *>(base + o_lparent);
    b.bounds = reinterpret_cast<float *>(base + o_bounds);
    b.stats = reinterpret_cast<uint32_t *>(base + o_stats);
    void *cub_tmp = base + o_cub;
    ctx->build_ids = b.ids_sorted;

    const unsigned T = 256, blocks = unsigned((nn + T - 1) / T);
    init_bounds_kernel<<<1, 32, 0, s>>>(reinterpret_cast<int *>(b.bounds));
    const float *boxes = d_boxes_in;
    if (d_vertices)
    {
        triangle_boxes_kernel<<<blocks, T, 0, s>>>(d_vertices, d_indices, n, b.boxes, reinterpret_cast<int *>(b.bounds));
        boxes = b.boxes;
    }
    else
    {
        box_bounds_kernel<<<blocks, T, 0, s>>>(boxes, n, reinterpret_cast<int *>(b.bounds));
    }
    morton_kernel<<<blocks, T, 0, s>>>(boxes, n, reinterpret_cast<const int *>(b.bounds), b.codes, b.ids);
    SPK_CUDA_TRY(cudaGetLastError());
    size_t tmp = std::max(sort_bytes, scan_bytes);
    SPK_CUDA_TRY(cub::DeviceRadixSort::SortPairs(cub_tmp, tmp, b.codes, b.codes_sorted, b.ids, b.ids_sorted, (int)n, 0, 30, s));
    ctx->stats.kernel_launches += 4;
    uint32_t stats_host[4] = {0, 0, 0, 0};
    if (n <= max_leaf)
    {
        // the whole mesh is one leaf: node 0 is a leaf, there are no pairs (but the root box is still wanted)
        out->num_pairs = 0;
        out->root_word = (leaf_base << 4) | n;
        out->depth = 0;
        std::vector<float> hb(size_t(n) * 6);
        SPK_CUDA_TRY(cudaMemcpyAsync(hb.data(), boxes, hb.size() * 4, cudaMemcpyDeviceToHost, s));
        SPK_CUDA_TRY(cudaStreamSynchronize(s));
        for (int a = 0; a < 3; a++)
        {
            out->root_bounds[a] = FLT_MAX;
            out->root_bounds[3 + a] = -FLT_MAX;
        }
        for (uint32_t k = 0; k < n; k++)
        {
            for (int a = 0; a < 3; a++)
            {
                out->root_bounds[a] = std::min(out->root_bounds[a], hb[size_t(k) * 6 + a]);
                out->root_bounds[3 + a] = std::max(out->root_bounds[3 + a], hb[size_t(k) * 6 + 3 + a]);
            }
        }
        out->sah_cost = double(n);
    }
    else
    {
        const unsigned iblocks = unsigned((nn - 1 + T - 1) / T);
        SPK_CUDA_TRY(cudaMemsetAsync(b.visits, 0, nn * 4, s));
        SPK_CUDA_TRY(cudaMemsetAsync(b.stats, 0, 64, s));
        radix_tree_kernel<<<iblocks, T, 0, s>>>(b.codes_sorted, (int)n, b.left, b.right, b.parent, b.leaf_parent, b.first, b.last);
        fit_bounds_kernel<<<blocks, T, 0, s>>>(boxes, b.ids_sorted, (int)n, b.left, b.right, b.parent, b.leaf_parent,
                                               b.node_boxes, b.visits);
        mark_inner_kernel<<<iblocks, T, 0, s>>>(b.first, b.last, (int)n, max_leaf, b.inner);
        SPK_CUDA_TRY(cudaGetLastError());
        // (the Morton codes are not needed any more: their buffer takes the prefix sum, one entry more than inner nodes)
        SPK_CUDA_TRY(cudaMemsetAsync(b.inner + (n - 1), 0, 4, s));
        tmp = std::max(sort_bytes, scan_bytes);
        SPK_CUDA_TRY(cub::DeviceScan::ExclusiveSum(cub_tmp, tmp, b.inner, b.codes, (int)n, s));
        preorder_slots_kernel<<<iblocks, T, 0, s>>>(b.left, b.parent, b.first, b.last, b.inner, b.codes, (int)n, b.slot, b.stats);
        uint32_t inner_count = 0;
        SPK_CUDA_TRY(cudaMemcpyAsync(&inner_count, b.codes + (n - 1), 4, cudaMemcpyDeviceToHost, s));
        SPK_CUDA_TRY(cudaStreamSynchronize(s));
        emit_pairs_kernel<<<iblocks, T, 0, s>>>(boxes, b.ids_sorted, (int)n, b.left, b.right, b.parent, b.first, b.last,
                                                b.node_boxes, b.inner, b.slot, pair_base, leaf_base, d_pairs, b.stats);
        SPK_CUDA_TRY(cudaGetLastError());
        ctx->stats.kernel_launches += 7;
        SPK_CUDA_TRY(cudaMemcpyAsync(stats_host, b.stats, 16, cudaMemcpyDeviceToHost, s));
        SPK_CUDA_TRY(cudaMemcpyAsync(out->root_bounds, b.node_boxes, 24, cudaMemcpyDeviceToHost, s));
        SPK_CUDA_TRY(cudaStreamSynchronize(s));
        out->num_pairs = inner_count;
        out->root_word = pair_base << 4; // the root is internal node 0 and holds slot 0
        out->depth = stats_host[0];
        const float *rb = out->root_bounds;
        const double root_area = double(rb[3] - rb[0]) * (rb[4] - rb[1]) + double(rb[4] - rb[1]) * (rb[5] - rb[2]) +
                                 double(rb[5] - rb[2]) * (rb[3] - rb[0]);
        float acc;
        memcpy(&acc, &stats_host[2], 4);
        out->sah_cost = root_area > 0 ? double(acc) / root_area : 0.0;
        if (out->depth > 32)
        {
            set_error("device BVH build: the tree is deeper than the 32-entry traversal stack (clustered input): hand in a "
                      "host-built tree for this mesh");
            return SPK_ERR_UNSUPPORTED;
        }
    }
    if (d_vertices && d_tris)
    {
        emit_triangles_kernel<<<blocks, T, 0, s>>>(d_vertices, d_indices, b.ids_sorted, n, d_tris);
        SPK_CUDA_TRY(cudaGetLastError());
        ctx->stats.kernel_launches++;
    }
    return SPK_OK;
}
} // namespace spk
