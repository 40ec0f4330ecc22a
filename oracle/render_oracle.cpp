// render_oracle.cpp -- CPU restatement of docs/RENDER_SPEC.md.  TEST
// INFRASTRUCTURE ONLY: it is linked by nothing in simpleraytracer_b200/ and may
// be executed only by tests/, __graft_entry__.smoke() and bench.py's CPU-baseline
// legs.
//
// PARITY UNPINNED: the reference (achernigin1987/SimpleRayTracer) contains no
// renderer, no tests and no golden vectors (SURVEY.md sections 0, 4, 8c), so
// there is nothing of the reference's to pin this oracle against.  It is a
// BUILD-DEFINED restatement: written independently of the CUDA kernels, from
// the spec text, as plain scalar C++ (one ray at a time, brute force over all
// triangles).  The committed fixtures under tests/golden/ freeze its output.
//
// A second, optional path (use_bvh = 1) walks a median-split BVH so that the
// CPU baseline can be timed at sizes where brute force is out of reach; the
// tests check that it returns exactly what the brute-force path returns.
//
// Build: g++ -O2 -ffp-contract=off -mfma -shared -fPIC (see oracle/Makefile):
// no implicit contraction, fmaf() is the hardware FMA.
#include <algorithm>
#include <atomic>
#include <chrono>
#include <cmath>
#include <cstdint>
#include <cstring>
#include <limits>
#include <thread>
#include <vector>

extern "C" {

struct orc_camera
{
    float origin[3], forward[3], right[3], up[3];
    float half_w, half_h;
};

struct orc_light
{
    float position[3], intensity[3];
};

struct orc_params
{
    uint32_t width, height;      // full frame
    uint32_t x0, x1, y0, y1;     // crop window [x0,x1) x [y0,y1); outputs are crop-sized
    uint32_t spp, jitter, seed;
    uint32_t shade;              // 0 albedo, 1 lambert, 2 lambert + shadow ray
    float t_min, shadow_eps;
    float background[3];
};

} // extern "C"

namespace {

struct Vec { float x, y, z; };

// --- spec section 0 ----------------------------------------------------------------
inline float Dot(Vec a, Vec b) { return std::fmaf(a.z, b.z, std::fmaf(a.y, b.y, a.x * b.x)); }

inline Vec Cross(Vec a, Vec b)
{
    return Vec{std::fmaf(a.y, b.z, -(a.z * b.y)),
               std::fmaf(a.z, b.x, -(a.x * b.z)),
               std::fmaf(a.x, b.y, -(a.y * b.x))};
}

inline Vec Sub(Vec a, Vec b) { return Vec{a.x - b.x, a.y - b.y, a.z - b.z}; }
inline Vec Scale(Vec a, float s) { return Vec{a.x * s, a.y * s, a.z * s}; }
inline Vec Load(float const* p) { return Vec{p[0], p[1], p[2]}; }

// --- spec section 1 ----------------------------------------------------------------
inline uint32_t Hash(uint32_t pix, uint32_t s, uint32_t dim, uint32_t seed)
{
    uint32_t h = (pix * 0x9E3779B1u) ^ (s * 0x85EBCA77u + dim * 0xC2B2AE3Du + seed * 0x27D4EB2Fu + 0x165667B1u);
    h ^= h >> 16;
    h *= 0x7FEB352Du;
    h ^= h >> 15;
    h *= 0x846CA68Bu;
    h ^= h >> 16;
    return h;
}

inline float U01(uint32_t h) { return static_cast<float>(h >> 8) * 5.9604644775390625e-08f; /* 2^-24 */ }

struct Triangle { Vec v0, e1, e2; };

struct SceneData
{
    std::vector<Triangle> tris;
    float const* albedo = nullptr; // 3 per triangle or null (0.8 grey)
    // optional BVH (culling only)
    struct Node { float lo[3], hi[3]; int left, right, first, count; };
    std::vector<Node> nodes;
    std::vector<int> order;
};

// --- spec section 3 ----------------------------------------------------------------
// Returns true when the triangle is `inside`; t receives ts/det.
inline bool Intersect(Vec o, Vec d, Triangle const& tri, float& t)
{
    Vec p = Cross(d, tri.e2);
    float det = Dot(tri.e1, p);
    Vec tv = Sub(o, tri.v0);
    float us = Dot(tv, p);
    Vec q = Cross(tv, tri.e1);
    float vs = Dot(d, q);
    float ts = Dot(tri.e2, q);
    if (det < 0.0f)
    {
        det = -det;
        us = -us;
        vs = -vs;
        ts = -ts;
    }
    bool inside = (det != 0.0f) && (us >= 0.0f) && (vs >= 0.0f) && ((us + vs) <= det);
    if (!inside) return false;
    t = ts / det;
    return true;
}

struct HitRecord { float t; int id; };

inline void Consider(HitRecord& best, float t, int id, float t_min)
{
    if (t > t_min && (t < best.t || (t == best.t && id < best.id)))
    {
        best.t = t;
        best.id = id;
    }
}

// --- BVH used only to cull (spec section 6) ---------------------------------------
struct RaySlabs { double inv[3]; double org[3]; };

inline bool SlabHit(SceneData::Node const& n, RaySlabs const& r, double t_hi)
{
    // double precision + generous slack: may only visit too much, never too little
    double lo = 0.0, hi = t_hi;
    for (int a = 0; a < 3; a++)
    {
        double t0 = (static_cast<double>(n.lo[a]) - r.org[a]) * r.inv[a];
        double t1 = (static_cast<double>(n.hi[a]) - r.org[a]) * r.inv[a];
        if (t0 > t1) std::swap(t0, t1);
        if (std::isnan(t0) || std::isnan(t1)) continue;
        lo = std::max(lo, t0);
        hi = std::min(hi, t1);
    }
    return lo <= hi * (1.0 + 1e-6) + 1e-12;
}

void BuildNode(SceneData& sc, std::vector<float> const& lo, std::vector<float> const& hi, int node,
               int first, int count)
{
    SceneData::Node n{};
    for (int a = 0; a < 3; a++) { n.lo[a] = std::numeric_limits<float>::max(); n.hi[a] = -n.lo[a]; }
    float clo[3] = {n.lo[0], n.lo[1], n.lo[2]}, chi[3] = {n.hi[0], n.hi[1], n.hi[2]};
    for (int i = first; i < first + count; i++)
    {
        int id = sc.order[i];
        for (int a = 0; a < 3; a++)
        {
            n.lo[a] = std::min(n.lo[a], lo[3 * id + a]);
            n.hi[a] = std::max(n.hi[a], hi[3 * id + a]);
            float c = 0.5f * (lo[3 * id + a] + hi[3 * id + a]);
            clo[a] = std::min(clo[a], c);
            chi[a] = std::max(chi[a], c);
        }
    }
    float pad = 0.f;
    for (int a = 0; a < 3; a++)
        pad = std::max(pad, std::max(std::fabs(n.lo[a]), std::fabs(n.hi[a])));
    pad = pad * 1e-5f + 1e-30f;
    for (int a = 0; a < 3; a++) { n.lo[a] -= pad; n.hi[a] += pad; }
    n.left = n.right = -1;
    n.first = first;
    n.count = count;
    if (count > 4)
    {
        int axis = 0;
        for (int a = 1; a < 3; a++)
            if (chi[a] - clo[a] > chi[axis] - clo[axis]) axis = a;
        int mid = first + count / 2;
        std::nth_element(sc.order.begin() + first, sc.order.begin() + mid, sc.order.begin() + first + count,
                         [&](int x, int y) {
                             float cx = lo[3 * x + axis] + hi[3 * x + axis];
                             float cy = lo[3 * y + axis] + hi[3 * y + axis];
                             return cx < cy || (cx == cy && x < y);
                         });
        n.left = static_cast<int>(sc.nodes.size());
        sc.nodes.emplace_back();
        n.right = static_cast<int>(sc.nodes.size());
        sc.nodes.emplace_back();
        n.count = 0;
        sc.nodes[node] = n;
        BuildNode(sc, lo, hi, n.left, first, mid - first);
        BuildNode(sc, lo, hi, n.right, mid, first + count - mid);
        return;
    }
    sc.nodes[node] = n;
}

void BuildBvh(SceneData& sc, float const* vertices)
{
    size_t n = sc.tris.size();
    std::vector<float> lo(3 * n), hi(3 * n);
    sc.order.resize(n);
    for (size_t i = 0; i < n; i++)
    {
        sc.order[i] = static_cast<int>(i);
        for (int a = 0; a < 3; a++)
        {
            float v0 = vertices[9 * i + a], v1 = vertices[9 * i + 3 + a], v2 = vertices[9 * i + 6 + a];
            lo[3 * i + a] = std::min(v0, std::min(v1, v2));
            hi[3 * i + a] = std::max(v0, std::max(v1, v2));
        }
    }
    sc.nodes.clear();
    sc.nodes.reserve(n + 16);
    sc.nodes.emplace_back();
    BuildNode(sc, lo, hi, 0, 0, static_cast<int>(n));
}

inline RaySlabs MakeSlabs(Vec o, Vec d)
{
    RaySlabs r;
    double dd[3] = {d.x, d.y, d.z}, oo[3] = {o.x, o.y, o.z};
    for (int a = 0; a < 3; a++)
    {
        r.inv[a] = 1.0 / dd[a]; // +-inf for zero components; NaNs are skipped in SlabHit
        r.org[a] = oo[a];
    }
    return r;
}

HitRecord ClosestHit(SceneData const& sc, Vec o, Vec d, float t_min, bool use_bvh)
{
    HitRecord best{std::numeric_limits<float>::infinity(), -1};
    if (!use_bvh)
    {
        for (size_t i = 0; i < sc.tris.size(); i++)
        {
            float t;
            if (Intersect(o, d, sc.tris[i], t)) Consider(best, t, static_cast<int>(i), t_min);
        }
        return best;
    }
    if (sc.tris.empty()) return best;
    RaySlabs r = MakeSlabs(o, d);
    std::vector<int> stack;
    stack.reserve(128);
    stack.push_back(0);
    while (!stack.empty())
    {
        SceneData::Node const& n = sc.nodes[stack.back()];
        stack.pop_back();
        double t_hi = std::isinf(best.t) ? std::numeric_limits<double>::infinity()
                                         : static_cast<double>(best.t) * (1.0 + 1e-4);
        if (!SlabHit(n, r, t_hi)) continue;
        if (n.left < 0)
        {
            for (int i = n.first; i < n.first + n.count; i++)
            {
                int id = sc.order[i];
                float t;
                if (Intersect(o, d, sc.tris[id], t)) Consider(best, t, id, t_min);
            }
        }
        else
        {
            stack.push_back(n.right);
            stack.push_back(n.left);
        }
    }
    return best;
}

bool AnyHit(SceneData const& sc, Vec o, Vec d, bool use_bvh)
{
    auto occludes = [&](Triangle const& tri) {
        float t;
        return Intersect(o, d, tri, t) && t > 0.0f && t < 1.0f;
    };
    if (!use_bvh)
    {
        for (auto const& tri : sc.tris)
            if (occludes(tri)) return true;
        return false;
    }
    if (sc.tris.empty()) return false;
    RaySlabs r = MakeSlabs(o, d);
    std::vector<int> stack;
    stack.reserve(128);
    stack.push_back(0);
    while (!stack.empty())
    {
        SceneData::Node const& n = sc.nodes[stack.back()];
        stack.pop_back();
        if (!SlabHit(n, r, 1.0 + 1e-4)) continue;
        if (n.left < 0)
        {
            for (int i = n.first; i < n.first + n.count; i++)
                if (occludes(sc.tris[sc.order[i]])) return true;
        }
        else
        {
            stack.push_back(n.right);
            stack.push_back(n.left);
        }
    }
    return false;
}

// --- spec section 2 ----------------------------------------------------------------
Vec PrimaryDirection(orc_camera const& cam, orc_params const& p, uint32_t x, uint32_t y, uint32_t s,
                     float inv2w, float inv2h)
{
    float jx = 0.5f, jy = 0.5f;
    if (p.jitter)
    {
        uint32_t pix = y * p.width + x;
        jx = U01(Hash(pix, s, 0u, p.seed));
        jy = U01(Hash(pix, s, 1u, p.seed));
    }
    float fx = static_cast<float>(x) + jx;
    float fy = static_cast<float>(y) + jy;
    float sx = std::fmaf(fx, inv2w, -1.0f) * cam.half_w;
    float sy = (-std::fmaf(fy, inv2h, -1.0f)) * cam.half_h;
    Vec r;
    r.x = std::fmaf(sy, cam.up[0], std::fmaf(sx, cam.right[0], cam.forward[0]));
    r.y = std::fmaf(sy, cam.up[1], std::fmaf(sx, cam.right[1], cam.forward[1]));
    r.z = std::fmaf(sy, cam.up[2], std::fmaf(sx, cam.right[2], cam.forward[2]));
    float len = std::sqrt(Dot(r, r));
    float inv = 1.0f / len;
    return Scale(r, inv);
}

// --- spec section 4 ----------------------------------------------------------------
Vec Shade(SceneData const& sc, orc_light const& light, orc_params const& p, Vec o, Vec d,
          HitRecord hit, bool use_bvh, uint64_t& shadow_rays)
{
    if (hit.id < 0) return Load(p.background);
    Vec alb = sc.albedo ? Load(sc.albedo + 3 * static_cast<size_t>(hit.id)) : Vec{0.8f, 0.8f, 0.8f};
    if (p.shade == 0) return alb;

    Triangle const& tri = sc.tris[static_cast<size_t>(hit.id)];
    Vec P{std::fmaf(hit.t, d.x, o.x), std::fmaf(hit.t, d.y, o.y), std::fmaf(hit.t, d.z, o.z)};
    Vec nr = Cross(tri.e1, tri.e2);
    float nl = std::sqrt(Dot(nr, nr));
    Vec n = Scale(nr, 1.0f / nl);
    if (Dot(n, d) > 0.0f) n = Vec{-n.x, -n.y, -n.z};
    Vec L = Sub(Load(light.position), P);
    float dist2 = Dot(L, L);
    float dist = std::sqrt(dist2);
    Vec l = Scale(L, 1.0f / dist);
    float c = Dot(n, l);
    if (!(c > 0.0f)) return Vec{0.f, 0.f, 0.f};
    if (p.shade == 2)
    {
        Vec Ps{std::fmaf(p.shadow_eps, n.x, P.x), std::fmaf(p.shadow_eps, n.y, P.y),
               std::fmaf(p.shadow_eps, n.z, P.z)};
        Vec ds = Sub(Load(light.position), Ps);
        shadow_rays++;
        if (AnyHit(sc, Ps, ds, use_bvh)) return Vec{0.f, 0.f, 0.f};
    }
    float k = c / dist2;
    return Vec{(alb.x * light.intensity[0]) * k, (alb.y * light.intensity[1]) * k,
               (alb.z * light.intensity[2]) * k};
}

} // namespace

double g_last_setup_ms = 0.0;

extern "C" {

// Host wall time the last orc_render call spent building its BVH (0 without use_bvh).
double orc_last_setup_ms() { return g_last_setup_ms; }

// Renders the crop window.  rgb: (y1-y0)*(x1-x0)*3 floats, hit_id/hit_t:
// (y1-y0)*(x1-x0); any may be NULL.  ray_counts[0] = primary, [1] = shadow.
// threads <= 0: all hardware threads.  Returns the number of threads used.
int orc_render(uint64_t n_tris, float const* vertices, float const* albedo, orc_camera const* cam,
               orc_light const* light, orc_params const* p, float* rgb, int32_t* hit_id, float* hit_t,
               uint64_t* ray_counts, int threads, int use_bvh)
{
    SceneData sc;
    sc.albedo = albedo;
    sc.tris.resize(n_tris);
    for (uint64_t i = 0; i < n_tris; i++)
    {
        Vec v0 = Load(vertices + 9 * i), v1 = Load(vertices + 9 * i + 3), v2 = Load(vertices + 9 * i + 6);
        sc.tris[i] = Triangle{v0, Sub(v1, v0), Sub(v2, v0)};
    }
    // scene preparation (edge precomputation above + the timing path's BVH) is timed separately: the GPU arm
    // reports its BVH build outside the frame time too (orc_last_setup_ms)
    auto const t_setup0 = std::chrono::steady_clock::now();
    if (use_bvh) BuildBvh(sc, vertices);
    g_last_setup_ms = std::chrono::duration<double, std::milli>(std::chrono::steady_clock::now() - t_setup0).count();

    float const inv2w = 2.0f / static_cast<float>(p->width);
    float const inv2h = 2.0f / static_cast<float>(p->height);
    float const inv_spp = 1.0f / static_cast<float>(p->spp);
    uint32_t const cw = p->x1 - p->x0, ch = p->y1 - p->y0;

    int n_threads = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    if (n_threads < 1) n_threads = 1;
    if (static_cast<uint32_t>(n_threads) > ch && ch > 0) n_threads = static_cast<int>(ch);

    std::atomic<uint32_t> next_row{0};
    std::atomic<uint64_t> total_primary{0}, total_shadow{0};
    auto worker = [&]() {
        uint64_t primary = 0, shadow = 0;
        Vec const o = Load(cam->origin);
        for (;;)
        {
            uint32_t row = next_row.fetch_add(1);
            if (row >= ch) break;
            uint32_t y = p->y0 + row;
            for (uint32_t col = 0; col < cw; col++)
            {
                uint32_t x = p->x0 + col;
                Vec acc{0.f, 0.f, 0.f};
                HitRecord first{std::numeric_limits<float>::infinity(), -1};
                for (uint32_t s = 0; s < p->spp; s++)
                {
                    Vec d = PrimaryDirection(*cam, *p, x, y, s, inv2w, inv2h);
                    primary++;
                    HitRecord hit = ClosestHit(sc, o, d, p->t_min, use_bvh != 0);
                    if (s == 0) first = hit;
                    Vec c = Shade(sc, *light, *p, o, d, hit, use_bvh != 0, shadow);
                    acc.x = acc.x + c.x;
                    acc.y = acc.y + c.y;
                    acc.z = acc.z + c.z;
                }
                size_t out = static_cast<size_t>(row) * cw + col;
                if (rgb)
                {
                    rgb[3 * out + 0] = acc.x * inv_spp;
                    rgb[3 * out + 1] = acc.y * inv_spp;
                    rgb[3 * out + 2] = acc.z * inv_spp;
                }
                if (hit_id) hit_id[out] = first.id;
                if (hit_t) hit_t[out] = first.t;
            }
        }
        total_primary += primary;
        total_shadow += shadow;
    };

    std::vector<std::thread> pool;
    for (int t = 1; t < n_threads; t++) pool.emplace_back(worker);
    worker();
    for (auto& th : pool) th.join();
    if (ray_counts)
    {
        ray_counts[0] = total_primary.load();
        ray_counts[1] = total_shadow.load();
    }
    return n_threads;
}

// Single-triangle probe for the degenerate-case tests: returns 1 when `inside`
// and writes t.
int orc_intersect(float const* o, float const* d, float const* v0v1v2, float* t)
{
    Vec v0 = Load(v0v1v2), v1 = Load(v0v1v2 + 3), v2 = Load(v0v1v2 + 6);
    Triangle tri{v0, Sub(v1, v0), Sub(v2, v0)};
    float tt = 0.f;
    bool inside = Intersect(Load(o), Load(d), tri, tt);
    if (inside && t) *t = tt;
    return inside ? 1 : 0;
}

} // extern "C"
