/* bvh_builder.cpp — see bvh_builder.h. Host-side product code (no oracle, no reference sources). */
#include "bvh_builder.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstring>
#include <vector>

namespace
{
struct Box
{
    float mn[3], mx[3];
    void Reset()
    {
        for (int a = 0; a < 3; a++)
        {
            mn[a] = FLT_MAX;
            mx[a] = -FLT_MAX;
        }
    }
    void Grow(const float *lo, const float *hi)
    {
        for (int a = 0; a < 3; a++)
        {
            mn[a] = std::min(mn[a], lo[a]);
            mx[a] = std::max(mx[a], hi[a]);
        }
    }
    void Grow(const Box &o)
    {
        Grow(o.mn, o.mx);
    }
    void GrowPoint(const float *p)
    {
        Grow(p, p);
    }
    float HalfArea() const
    {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        if (dx < 0.f || dy < 0.f || dz < 0.f)
        {
            return 0.f;
        }
        return dx * dy + dy * dz + dz * dx;
    }
};

struct NodeRec
{
    float bounds[6];
    uint32_t index;
    uint32_t pad;
};
static_assert(sizeof(NodeRec) == 32, "SpkBvhNode layout");

constexpr int kBins = 16;

struct Builder
{
    const float *bboxes;
    const float *centers;
    uint32_t max_leaf;
    uint32_t max_depth;
    std::vector<NodeRec> nodes;
    std::vector<uint32_t> ids;
    uint32_t depth_seen = 0;

    static void SetBounds(NodeRec &n, const Box &b)
    {
        for (int a = 0; a < 3; a++)
        {
            n.bounds[2 * a] = b.mn[a];
            n.bounds[2 * a + 1] = b.mx[a];
        }
    }

    Box RangeBox(uint32_t begin, uint32_t end) const
    {
        Box b;
        b.Reset();
        for (uint32_t k = begin; k < end; k++)
        {
            const float *p = bboxes + size_t(ids[k]) * 6;
            b.Grow(p, p + 3);
        }
        return b;
    }

    // smallest d with max_leaf * 2^d >= count
    uint32_t LevelsNeeded(uint32_t count) const
    {
        uint32_t d = 0;
        uint64_t cap = max_leaf;
        while (cap < count)
        {
            cap *= 2;
            d++;
        }
        return d;
    }

    void Build(uint32_t node_id, uint32_t begin, uint32_t end, uint32_t depth)
    {
        depth_seen = std::max(depth_seen, depth);
        uint32_t count = end - begin;
        Box cb;
        cb.Reset();
        for (uint32_t k = begin; k < end; k++)
        {
            cb.GrowPoint(centers + size_t(ids[k]) * 3);
        }
        const Box node_box = RangeBox(begin, end);
        SetBounds(nodes[node_id], node_box);

        bool must_split = count > max_leaf;
        bool budget_tight = depth + LevelsNeeded(count) + 1 >= max_depth;
        uint32_t mid = 0;
        bool split = false;

        if (count > 1 && !budget_tight)
        {
            // binned SAH over the three axes
            float best_cost = FLT_MAX;
            int best_axis = -1;
            int best_bin = -1;
            for (int axis = 0; axis < 3; axis++)
            {
                float lo = cb.mn[axis], hi = cb.mx[axis];
                if (!(hi > lo))
                {
                    continue;
                }
                Box bin_box[kBins];
                uint32_t bin_count[kBins];
                for (int b = 0; b < kBins; b++)
                {
                    bin_box[b].Reset();
                    bin_count[b] = 0;
                }
                float scale = float(kBins) / (hi - lo);
                for (uint32_t k = begin; k < end; k++)
                {
                    uint32_t id = ids[k];
                    int b = std::min(kBins - 1, std::max(0, int((centers[size_t(id) * 3 + axis] - lo) * scale)));
                    bin_count[b]++;
                    bin_box[b].Grow(bboxes + size_t(id) * 6, bboxes + size_t(id) * 6 + 3);
                }
                float right_area[kBins];
                uint32_t right_count[kBins];
                Box acc;
                acc.Reset();
                uint32_t n = 0;
                for (int b = kBins - 1; b > 0; b--)
                {
                    acc.Grow(bin_box[b]);
                    n += bin_count[b];
                    right_area[b] = acc.HalfArea();
                    right_count[b] = n;
                }
                acc.Reset();
                n = 0;
                for (int b = 0; b < kBins - 1; b++)
                {
                    acc.Grow(bin_box[b]);
                    n += bin_count[b];
                    if (n == 0 || right_count[b + 1] == 0)
                    {
                        continue;
                    }
                    float cost = acc.HalfArea() * float(n) + right_area[b + 1] * float(right_count[b + 1]);
                    if (cost < best_cost)
                    {
                        best_cost = cost;
                        best_axis = axis;
                        best_bin = b;
                    }
                }
            }
            if (best_axis >= 0)
            {
                float parent_area = node_box.HalfArea();
                float leaf_cost = float(count);
                float split_cost = 1.0f + (parent_area > 0.f ? best_cost / parent_area : FLT_MAX);
                if (must_split || split_cost < leaf_cost)
                {
                    float lo = cb.mn[best_axis], hi = cb.mx[best_axis];
                    float scale = float(kBins) / (hi - lo);
                    auto it = std::partition(ids.begin() + begin, ids.begin() + end, [&](uint32_t id) {
                        int b = std::min(kBins - 1,
                                         std::max(0, int((centers[size_t(id) * 3 + best_axis] - lo) * scale)));
                        return b <= best_bin;
                    });
                    mid = uint32_t(it - ids.begin());
                    split = mid > begin && mid < end;
                }
            }
        }
        if (!split && (must_split || (budget_tight && count > max_leaf)))
        {
            // object median on the widest centroid axis (also the depth-budget fallback)
            int axis = 0;
            float ext = -1.f;
            for (int a = 0; a < 3; a++)
            {
                float e = cb.mx[a] - cb.mn[a];
                if (e > ext)
                {
                    ext = e;
                    axis = a;
                }
            }
            mid = begin + count / 2;
            std::nth_element(ids.begin() + begin, ids.begin() + mid, ids.begin() + end, [&](uint32_t x, uint32_t y) {
                float cx = centers[size_t(x) * 3 + axis], cy = centers[size_t(y) * 3 + axis];
                return cx < cy || (cx == cy && x < y);
            });
            split = true;
        }
        if (!split)
        {
            nodes[node_id].index = (begin << 4) | count;
            return;
        }
        uint32_t first_child = uint32_t(nodes.size());
        nodes.push_back(NodeRec{});
        nodes.push_back(NodeRec{});
        nodes[node_id].index = first_child << 4;
        Build(first_child, begin, mid, depth + 1);
        Build(first_child + 1, mid, end, depth + 1);
    }
};
} // namespace

struct SpkHostBvh
{
    std::vector<NodeRec> nodes;
    std::vector<uint32_t> prim_ids;
    uint32_t depth = 0;
};

extern "C"
{
SpkHostBvh *spk_host_bvh_build(const float *bboxes, const float *centers, uint32_t num_prims, uint32_t max_leaf,
                               uint32_t max_depth)
{
    if (num_prims == 0 || !bboxes || max_leaf < 1 || max_leaf > 15 || max_depth < 2 || max_depth > 32 ||
        num_prims >= (1u << 28))
    {
        return nullptr;
    }
    std::vector<float> own_centers;
    if (!centers)
    {
        own_centers.resize(size_t(num_prims) * 3);
        for (uint32_t i = 0; i < num_prims; i++)
        {
            for (int a = 0; a < 3; a++)
            {
                own_centers[size_t(i) * 3 + a] = (bboxes[size_t(i) * 6 + a] + bboxes[size_t(i) * 6 + 3 + a]) / 2.f;
            }
        }
        centers = own_centers.data();
    }
    Builder b;
    b.bboxes = bboxes;
    b.centers = centers;
    b.max_leaf = max_leaf;
    b.max_depth = max_depth;
    b.ids.resize(num_prims);
    for (uint32_t i = 0; i < num_prims; i++)
    {
        b.ids[i] = i;
    }
    // a full tree of `max_leaf`-sized leaves must fit the depth budget
    if (b.LevelsNeeded(num_prims) + 1 >= max_depth)
    {
        return nullptr;
    }
    b.nodes.reserve(size_t(num_prims) / std::max(1u, max_leaf / 2) + 16);
    b.nodes.push_back(NodeRec{});
    b.Build(0, 0, num_prims, 0);
    auto *out = new SpkHostBvh;
    out->nodes = std::move(b.nodes);
    out->prim_ids = std::move(b.ids);
    out->depth = b.depth_seen;
    return out;
}

SpkHostBvh *spk_host_bvh_build_triangles(const float *vertices, const uint32_t *indices, uint32_t num_faces,
                                         uint32_t max_leaf, uint32_t max_depth)
{
    if (!vertices || !indices || num_faces == 0)
    {
        return nullptr;
    }
    std::vector<float> boxes(size_t(num_faces) * 6);
    std::vector<float> centers(size_t(num_faces) * 3);
    for (uint32_t f = 0; f < num_faces; f++)
    {
        float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
        float mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int k = 0; k < 3; k++)
        {
            const float *v = vertices + size_t(indices[size_t(f) * 3 + k]) * 3;
            for (int a = 0; a < 3; a++)
            {
                mn[a] = std::min(mn[a], v[a]);
                mx[a] = std::max(mx[a], v[a]);
            }
        }
        for (int a = 0; a < 3; a++)
        {
            boxes[size_t(f) * 6 + a] = mn[a];
            boxes[size_t(f) * 6 + 3 + a] = mx[a];
            centers[size_t(f) * 3 + a] = (mn[a] + mx[a]) / 2.f;
        }
    }
    return spk_host_bvh_build(boxes.data(), centers.data(), num_faces, max_leaf, max_depth);
}

void spk_host_bvh_free(SpkHostBvh *bvh)
{
    delete bvh;
}
uint32_t spk_host_bvh_num_nodes(const SpkHostBvh *bvh)
{
    return uint32_t(bvh->nodes.size());
}
uint32_t spk_host_bvh_num_prims(const SpkHostBvh *bvh)
{
    return uint32_t(bvh->prim_ids.size());
}
uint32_t spk_host_bvh_depth(const SpkHostBvh *bvh)
{
    return bvh->depth;
}
const void *spk_host_bvh_nodes(const SpkHostBvh *bvh)
{
    return bvh->nodes.data();
}
const uint32_t *spk_host_bvh_prim_ids(const SpkHostBvh *bvh)
{
    return bvh->prim_ids.data();
}
double spk_host_bvh_sah_cost(const SpkHostBvh *bvh)
{
    const auto &n = bvh->nodes;
    auto area = [](const NodeRec &r) {
        double dx = r.bounds[1] - r.bounds[0], dy = r.bounds[3] - r.bounds[2], dz = r.bounds[5] - r.bounds[4];
        return dx * dy + dy * dz + dz * dx;
    };
    double root = area(n[0]);
    if (!(root > 0.0))
    {
        return 0.0;
    }
    double cost = 0.0;
    for (const auto &r : n)
    {
        uint32_t count = r.index & 15u;
        cost += area(r) * (count ? double(count) : 1.0);
    }
    return cost / root;
}
}
