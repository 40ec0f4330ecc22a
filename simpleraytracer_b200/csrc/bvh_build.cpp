// bvh_build.cpp -- deterministic host-side BVH construction (SURVEY.md row R3;
// BUILD-DEFINED: the reference has no acceleration structure).
//
// Top-down binned SAH (16 bins x 3 axes), leaves of <= kMaxLeafTris triangles,
// balanced median fallback past depth 40 so that the traversal stack
// (kBvhStackDepth) can never overflow.  The tree only culls (docs/RENDER_SPEC.md
// section 6): child boxes are padded by 2^-18 of the scene diagonal so rounding in the
// slab test can never drop a triangle the Moller-Trumbore test would accept.
#include "scene.h"

#include <algorithm>
#include <cfloat>
#include <cmath>
#include <cstdlib>

namespace ML {
namespace {

struct Prim
{
    float bmin[3], bmax[3], c[3];
    int32_t id;
};

struct Box
{
    float mn[3] = {FLT_MAX, FLT_MAX, FLT_MAX};
    float mx[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    void Grow(float const* lo, float const* hi)
    {
        for (int a = 0; a < 3; a++)
        {
            mn[a] = std::min(mn[a], lo[a]);
            mx[a] = std::max(mx[a], hi[a]);
        }
    }
    void Grow(Box const& b) { Grow(b.mn, b.mx); }
    float HalfArea() const
    {
        float dx = mx[0] - mn[0], dy = mx[1] - mn[1], dz = mx[2] - mn[2];
        if (dx < 0.f) return 0.f;
        return dx * dy + dy * dz + dz * dx;
    }
};

constexpr int kMaxBins = 64;
static int const kBins = [] {   // SAH bins per axis (SRT_BVH_BINS=4..64, experiments)
    char const* e = std::getenv("SRT_BVH_BINS");
    int v = e ? std::atoi(e) : 16;
    return v < 4 ? 4 : (v > kMaxBins ? kMaxBins : v);
}();
constexpr int kMedianDepth = 40;

inline float IntBitsToFloat(int32_t v)
{
    float f;
    std::memcpy(&f, &v, 4);
    return f;
}

class Builder
{
public:
    Builder(size_t n, float const* vertices) : m_prims(n)
    {
        Box scene;
        for (size_t i = 0; i < n; i++)
        {
            Prim& p = m_prims[i];
            float const* v = vertices + 9 * i;
            for (int a = 0; a < 3; a++)
            {
                p.bmin[a] = std::min(v[a], std::min(v[3 + a], v[6 + a]));
                p.bmax[a] = std::max(v[a], std::max(v[3 + a], v[6 + a]));
                p.c[a] = 0.5f * (p.bmin[a] + p.bmax[a]);
            }
            p.id = static_cast<int32_t>(i);
            scene.Grow(p.bmin, p.bmax);
        }
        if (n > 0)
        {
            float dx = scene.mx[0] - scene.mn[0], dy = scene.mx[1] - scene.mn[1],
                  dz = scene.mx[2] - scene.mn[2];
            float diag = std::sqrt(dx * dx + dy * dy + dz * dz);
            float mag = 0.f;
            for (int a = 0; a < 3; a++)
                mag = std::max(mag, std::max(std::fabs(scene.mn[a]), std::fabs(scene.mx[a])));
            m_pad = std::max(diag, mag) * (1.0f / 262144.0f) + FLT_MIN;
        }
    }

    BvhBuildResult Run()
    {
        BvhBuildResult out;
        size_t const n = m_prims.size();
        m_nodes.reserve(n / 2 + 2);
        if (n <= static_cast<size_t>(m_leaf_size))
        {
            // Root must be an inner node: both children reference the one leaf
            // (for n == 0: the first padding triangle, which can never be hit).
            Box b;
            for (auto const& p : m_prims) b.Grow(p.bmin, p.bmax);
            if (n == 0) { b.mn[0] = b.mn[1] = b.mn[2] = 0.f; b.mx[0] = b.mx[1] = b.mx[2] = 0.f; }
            int32_t ref = LeafRef(0, n == 0 ? 1 : static_cast<int>(n));
            m_nodes.emplace_back();
            StoreNode(0, ref, b, ref, b);
            m_max_depth = 1;
        }
        else
        {
            Box b;
            BuildRange(0, static_cast<int>(n), 0, b);
        }
        out.nodes = std::move(m_nodes);
        out.order.resize(n);
        for (size_t i = 0; i < n; i++) out.order[i] = m_prims[i].id;
        out.max_depth = m_max_depth;
        if (m_max_depth >= kBvhStackDepth - 2)
            throw std::runtime_error("BVH depth exceeds traversal stack");
        return out;
    }

private:
    static int32_t LeafRef(int first, int count)
    {
        return ~((first << kLeafCountBits) | (count - 1));
    }

    void StoreNode(int idx, int32_t ref0, Box const& b0, int32_t ref1, Box const& b1)
    {
        float const p = m_pad;
        BvhNode& n = m_nodes[idx];
        // per axis: (c0.min, c1.min, c0.max, c1.max) -- the two children side by
        // side, so one packed f32x2 FMA handles both boxes
        n.n0 = {b0.mn[0] - p, b1.mn[0] - p, b0.mx[0] + p, b1.mx[0] + p};
        n.n1 = {b0.mn[1] - p, b1.mn[1] - p, b0.mx[1] + p, b1.mx[1] + p};
        n.n2 = {b0.mn[2] - p, b1.mn[2] - p, b0.mx[2] + p, b1.mx[2] + p};
        n.n3 = {IntBitsToFloat(ref0), IntBitsToFloat(ref1), 0.f, 0.f};
    }

    // Builds the subtree over prims [first, first+count); returns its child
    // ref and bounds.
    int32_t BuildRange(int first, int count, int depth, Box& bounds)
    {
        m_max_depth = std::max(m_max_depth, depth);
        if (count <= m_leaf_size)
        {
            for (int i = first; i < first + count; i++) bounds.Grow(m_prims[i].bmin, m_prims[i].bmax);
            return LeafRef(first, count);
        }

        int mid = -1;
        if (depth < kMedianDepth) mid = SahSplit(first, count);
        if (mid < 0) mid = MedianSplit(first, count);

        int const idx = static_cast<int>(m_nodes.size());
        m_nodes.emplace_back();
        Box b0, b1;
        int32_t ref0 = BuildRange(first, mid - first, depth + 1, b0);
        int32_t ref1 = BuildRange(mid, first + count - mid, depth + 1, b1);
        StoreNode(idx, ref0, b0, ref1, b1);
        bounds.Grow(b0);
        bounds.Grow(b1);
        return idx;
    }

    int SahSplit(int first, int count)
    {
        float cmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, cmax[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
        for (int i = first; i < first + count; i++)
            for (int a = 0; a < 3; a++)
            {
                cmin[a] = std::min(cmin[a], m_prims[i].c[a]);
                cmax[a] = std::max(cmax[a], m_prims[i].c[a]);
            }

        float best_cost = FLT_MAX;
        int best_axis = -1, best_bin = -1;
        float best_scale = 0.f;
        for (int a = 0; a < 3; a++)
        {
            float extent = cmax[a] - cmin[a];
            if (!(extent > 0.f)) continue;
            float scale = static_cast<float>(kBins) / extent;
            Box bin_box[kMaxBins];
            int bin_cnt[kMaxBins] = {};
            for (int i = first; i < first + count; i++)
            {
                int b = BinOf(m_prims[i].c[a], cmin[a], scale);
                bin_cnt[b]++;
                bin_box[b].Grow(m_prims[i].bmin, m_prims[i].bmax);
            }
            float right_area[kMaxBins];
            int right_cnt[kMaxBins];
            Box acc;
            int cnt = 0;
            for (int b = kBins - 1; b > 0; b--)
            {
                acc.Grow(bin_box[b]);
                cnt += bin_cnt[b];
                right_area[b] = acc.HalfArea();
                right_cnt[b] = cnt;
            }
            Box lacc;
            int lcnt = 0;
            for (int b = 0; b < kBins - 1; b++)
            {
                lacc.Grow(bin_box[b]);
                lcnt += bin_cnt[b];
                if (lcnt == 0 || right_cnt[b + 1] == 0) continue;
                float cost = lacc.HalfArea() * static_cast<float>(lcnt)
                           + right_area[b + 1] * static_cast<float>(right_cnt[b + 1]);
                if (cost < best_cost)
                {
                    best_cost = cost;
                    best_axis = a;
                    best_bin = b;
                    best_scale = scale;
                }
            }
        }
        if (best_axis < 0) return -1;

        int const a = best_axis;
        float const lo = cmin[a];
        auto it = std::partition(m_prims.begin() + first, m_prims.begin() + first + count,
                                 [=](Prim const& p) { return BinOf(p.c[a], lo, best_scale) <= best_bin; });
        int mid = static_cast<int>(it - m_prims.begin());
        if (mid == first || mid == first + count) return -1;
        return mid;
    }

    int MedianSplit(int first, int count)
    {
        Box cb;
        for (int i = first; i < first + count; i++) cb.Grow(m_prims[i].c, m_prims[i].c);
        int a = 0;
        float best = cb.mx[0] - cb.mn[0];
        for (int k = 1; k < 3; k++)
            if (cb.mx[k] - cb.mn[k] > best) { best = cb.mx[k] - cb.mn[k]; a = k; }
        int mid = first + count / 2;
        std::nth_element(m_prims.begin() + first, m_prims.begin() + mid,
                         m_prims.begin() + first + count, [a](Prim const& x, Prim const& y) {
                             return x.c[a] < y.c[a] || (x.c[a] == y.c[a] && x.id < y.id);
                         });
        return mid;
    }

    static int BinOf(float c, float lo, float scale)
    {
        int b = static_cast<int>((c - lo) * scale);
        return b < 0 ? 0 : (b >= kBins ? kBins - 1 : b);
    }

    std::vector<Prim> m_prims;
    std::vector<BvhNode> m_nodes;
    float m_pad = 0.f;
    int m_max_depth = 0;
    // tuning knob for experiments only (SRT_BVH_LEAF=1..kMaxLeafTris)
    int m_leaf_size = [] {
        char const* e = std::getenv("SRT_BVH_LEAF");
        int v = e ? std::atoi(e) : kDefaultLeafTris;
        return v < 1 ? 1 : (v > kMaxLeafTris ? kMaxLeafTris : v);
    }();
};

} // namespace

BvhBuildResult BuildBvh(size_t n, float const* vertices)
{
    if (n >= (size_t(1) << (31 - kLeafCountBits)))
        throw std::runtime_error("Too many triangles for the BVH leaf encoding");
    Builder b(n, vertices);
    return b.Run();
}

} // namespace ML
