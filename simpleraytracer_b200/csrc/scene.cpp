// scene.cpp -- scene object: SoA upload, BVH, render dispatch and the srt* C
// entry points (include/srt_render.h).  BUILD-DEFINED (no reference code);
// error conventions follow the reference's: factories report on the context
// (context.cpp:20-49), operations return ML_FAIL and leave a message on the
// object (model.cpp:161-176).
#include "scene.h"

#include <cstdio>
#include <cstdlib>

#include "context.h"
#include "image.h"

#include <chrono>
#include <cfloat>
#include <cmath>

namespace ML {

namespace {
inline float IntBitsToFloat(int32_t v)
{
    float f;
    std::memcpy(&f, &v, 4);
    return f;
}

// The 8-row blocks a call owns: block k = 0..count-1 starts at frame row
// origin + k*step (see RenderLaunch).  stride <= 1: the whole band from y0.
void RowBlocks(srt_render_params const& p, uint32_t& origin, uint32_t& step, uint32_t& count)
{
    uint32_t const rows = p.y1 - p.y0;
    uint32_t const blocks = (rows + 7u) / 8u;
    uint32_t const stride = p.row_block_stride > 1 ? p.row_block_stride : 1u;
    if (stride == 1)
    {
        origin = p.y0;
        step = 8;
        count = blocks;
        return;
    }
    // frame block index b = y/8 (y0 % 8 == 0 is validated); owned: b % stride == phase
    uint32_t const b0 = p.y0 / 8u;
    uint32_t const first = b0 + (p.row_block_phase + stride - b0 % stride) % stride;
    origin = first * 8u;
    step = 8u * stride;
    count = first < b0 + blocks ? (b0 + blocks - first + stride - 1) / stride : 0;
}

// Per-frame device scratch (one 256-byte memset per render call): 4 x u64 ray/test counters | 2 x u32 block
// tickets of the completion signal (one per launch stream) | pad | GridFlags of the screen grid | block trace.
constexpr size_t kScratchBytes = 256;
constexpr size_t kScratchTicketWord = 8;    // in uint32 units
constexpr size_t kScratchGridFlagsWord = 12;
constexpr size_t kScratchBucketWord = 32;    // kGridBuckets x u32 item counters of the work list
constexpr size_t kScratchWorkWord = 44;      // u32 cursor of the persistent render blocks
constexpr size_t kScratchTraceU64 = 8;      // 8 x u64 block timeline probe (SRT_BLOCK_TRACE=1): words 16..31, + entry / exit stamps in words 40..43
} // namespace

// Rows [y0, y1) cut into at most `requested` chunks of whole (8 * stride)-row groups, so that every chunk
// holds the same number of row blocks of every rank of an interleaved frame.
void ChunkLayout(srt_render_params const& p, uint32_t requested, uint32_t& chunk_rows, uint32_t& n_chunks)
{
    uint32_t const rows = p.y1 - p.y0;
    uint32_t const group = 8u * (p.row_block_stride > 1 ? p.row_block_stride : 1u);
    uint32_t const want = requested > 0 ? requested : 1u;
    chunk_rows = ((rows + want - 1u) / want + group - 1u) / group * group;
    if (chunk_rows == 0) chunk_rows = group;
    n_chunks = rows ? (rows + chunk_rows - 1u) / chunk_rows : 0u;
}

Scene::Scene(DeviceStatePtr device, size_t n, float const* vertices, float const* albedo,
             bool build_bvh, BvhBuildResult const* prebuilt)
    : m_device(std::move(device))
{
    if (n > 0 && vertices == nullptr) throw std::runtime_error("Bad vertices argument");
    if (n >= (size_t(1) << 27)) throw std::runtime_error("Too many triangles");
    for (size_t i = 0; i < 9 * n; i++)
        if (!std::isfinite(vertices[i])) throw std::runtime_error("Non-finite vertex coordinate");

    m_device->MakeCurrent();

    // ---- BVH (host) ---------------------------------------------------------
    std::vector<int32_t> order(n);
    for (size_t i = 0; i < n; i++) order[i] = static_cast<int32_t>(i);
    auto t0 = std::chrono::steady_clock::now();
    if (build_bvh)
    {
        // a scene group builds the BVH once and hands it to the replica of every device
        BvhBuildResult bvh = prebuilt ? *prebuilt : BuildBvh(n, vertices);
        m_nodes_host = std::move(bvh.nodes);
        order = bvh.order;
    }
    auto t1 = std::chrono::steady_clock::now();
    m_order_host = order;
    m_info.bvh_build_ms = std::chrono::duration<float, std::milli>(t1 - t0).count();

    // ---- SoA in leaf order, padded to the tile size ---------------------------
    size_t const padded = ((n + kTriTile - 1) / kTriTile) * kTriTile + (n == 0 ? kTriTile : 0);
    std::vector<F4> soa(3 * padded, F4{0.f, 0.f, 0.f, 0.f});
    F4* v0 = soa.data();
    F4* e1 = v0 + padded;
    F4* e2 = e1 + padded;
    for (size_t i = 0; i < padded; i++) v0[i].w = IntBitsToFloat(-1);
    float bmin[3] = {FLT_MAX, FLT_MAX, FLT_MAX}, bmax[3] = {-FLT_MAX, -FLT_MAX, -FLT_MAX};
    for (size_t i = 0; i < n; i++)
    {
        int32_t id = order[i];
        float const* v = vertices + 9 * static_cast<size_t>(id);
        v0[i] = {v[0], v[1], v[2], IntBitsToFloat(id)};
        e1[i] = {v[3] - v[0], v[4] - v[1], v[5] - v[2], 0.f};
        e2[i] = {v[6] - v[0], v[7] - v[1], v[8] - v[2], 0.f};
        for (int k = 0; k < 3; k++)
        {
            bmin[k] = std::min(bmin[k], std::min(v[k], std::min(v[3 + k], v[6 + k])));
            bmax[k] = std::max(bmax[k], std::max(v[k], std::max(v[3 + k], v[6 + k])));
        }
    }
    std::vector<F4> alb(std::max<size_t>(n, 1), F4{0.8f, 0.8f, 0.8f, 0.f});
    if (albedo)
        for (size_t i = 0; i < n; i++) alb[i] = {albedo[3 * i], albedo[3 * i + 1], albedo[3 * i + 2], 0.f};

    // ---- upload ----------------------------------------------------------------
    size_t const tri_bytes = soa.size() * sizeof(F4);
    size_t const nrm_bytes = padded * sizeof(F4);      // tri_n, behind the three uploaded arrays
    size_t const alb_bytes = alb.size() * sizeof(F4);
    size_t const node_bytes = m_nodes_host.size() * sizeof(BvhNode);
    try
    {
        ML_CUDA_CHECK(cudaMalloc(&m_d_tris, tri_bytes + nrm_bytes));
        ML_CUDA_CHECK(cudaMalloc(&m_d_albedo, alb_bytes));
        if (node_bytes) ML_CUDA_CHECK(cudaMalloc(&m_d_nodes, node_bytes));
        ML_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&m_d_counters), kScratchBytes));
        cudaStream_t s = m_device->stream;
        ML_CUDA_CHECK(cudaEventRecord(m_device->ev_start, s));
        ML_CUDA_CHECK(cudaMemcpyAsync(m_d_tris, soa.data(), tri_bytes, cudaMemcpyHostToDevice, s));
        {
            F4* const base = static_cast<F4*>(m_d_tris);
            LaunchTriNormals(base + padded, base + 2 * padded, base + 3 * padded, static_cast<uint32_t>(padded), s);
        }
        ML_CUDA_CHECK(cudaMemcpyAsync(m_d_albedo, alb.data(), alb_bytes, cudaMemcpyHostToDevice, s));
        if (node_bytes)
            ML_CUDA_CHECK(cudaMemcpyAsync(m_d_nodes, m_nodes_host.data(), node_bytes, cudaMemcpyHostToDevice, s));
        ML_CUDA_CHECK(cudaMemsetAsync(m_d_counters, 0, kScratchBytes, s));
        ML_CUDA_CHECK(cudaEventRecord(m_device->ev_stop, s));
        ML_CUDA_CHECK(cudaStreamSynchronize(s));
        ML_CUDA_CHECK(cudaEventElapsedTime(&m_info.upload_ms, m_device->ev_start, m_device->ev_stop));
    }
    catch (...)
    {
        cudaFree(m_d_tris); cudaFree(m_d_albedo); cudaFree(m_d_nodes); cudaFree(m_d_counters);
        throw;
    }

    m_view.tri_v0 = static_cast<F4 const*>(m_d_tris);
    m_view.tri_e1 = m_view.tri_v0 + padded;
    m_view.tri_e2 = m_view.tri_e1 + padded;
    m_view.tri_n = m_view.tri_e2 + padded;
    m_view.albedo = static_cast<F4 const*>(m_d_albedo);
    m_view.nodes = static_cast<F4 const*>(m_d_nodes);
    m_view.n_tris = static_cast<uint32_t>(n);
    m_view.n_tris_padded = static_cast<uint32_t>(padded);

    m_info.n_triangles = n;
    m_info.n_bvh_nodes = m_nodes_host.size();
    m_info.device_bytes = tri_bytes + nrm_bytes + alb_bytes + node_bytes;
    for (int k = 0; k < 3; k++)
    {
        m_info.bounds_min[k] = n ? bmin[k] : 0.f;
        m_info.bounds_max[k] = n ? bmax[k] : 0.f;
    }
    if (n > 0)
    {
        // same slack as the BVH boxes (bvh_build.cpp): 2^-18 of the scene extent
        float dx = bmax[0] - bmin[0], dy = bmax[1] - bmin[1], dz = bmax[2] - bmin[2];
        float diag = std::sqrt(dx * dx + dy * dy + dz * dz), mag = 0.f;
        for (int k = 0; k < 3; k++) mag = std::max(mag, std::max(std::fabs(bmin[k]), std::fabs(bmax[k])));
        m_pad = std::max(diag, mag) * (1.0f / 262144.0f) + FLT_MIN;
    }
}

void Scene::FreeGrid()
{
    cudaFree(m_d_grid_tiles); cudaFree(m_d_grid_items); cudaFree(m_d_grid_bbox); cudaFree(m_d_grid_misc);
    cudaFree(m_d_grid_order);
    m_d_grid_tiles = m_d_grid_items = m_d_grid_bbox = m_d_grid_misc = m_d_grid_order = nullptr;
    m_grid_tiles = m_grid_cap = 0;
    m_grid = GridView{};
    cudaFree(m_d_lgrid_tiles); cudaFree(m_d_lgrid_items); cudaFree(m_d_lgrid_bbox); cudaFree(m_d_lgrid_misc);
    m_d_lgrid_tiles = m_d_lgrid_items = m_d_lgrid_bbox = m_d_lgrid_misc = nullptr;
    m_lgrid_tiles = m_lgrid_cap = 0;
    m_lgrid = LightGridView{};
    m_lgrid_valid = false;
}

// Shadow-ray grid for this call's light (scene.h, LightGridView), built on `stream`.  Used when the
// whole scene lies in front of a projection plane through the light (a light above, beside or outside
// the geometry); otherwise m_lgrid.grid.enabled stays 0 and shadow rays walk the BVH.
void Scene::PrepareLightGrid(srt_light const* light, srt_render_params const* params, cudaStream_t stream)
{
    m_lgrid.grid.enabled = 0;
    int const want = std::getenv("SRT_LIGHT_GRID") ? std::atoi(std::getenv("SRT_LIGHT_GRID")) : 1;   // read per call: tests flip it
    if (!want || params->accel != SRT_ACCEL_BVH || !(RenderUsesGrid(*params) & 2) || m_view.n_tris == 0) return;
    // The light grid depends on the scene and the light position only (not on the camera or the frame):
    // it is kept while the light stays where it is, like the BVH is kept while the scene does.
    static int const cache = std::getenv("SRT_LIGHT_GRID_CACHE") ? std::atoi(std::getenv("SRT_LIGHT_GRID_CACHE")) : 1;
    if (cache && m_lgrid_valid && std::memcmp(m_lgrid_light, light->position, sizeof(m_lgrid_light)) == 0)
    {
        ML_CUDA_CHECK(cudaStreamWaitEvent(stream, m_lgrid_built, 0));   // built on another stream, perhaps
        m_lgrid.grid.enabled = 1;
        return;
    }
    m_lgrid_valid = false;
    double const lo[3] = {m_info.bounds_min[0], m_info.bounds_min[1], m_info.bounds_min[2]};
    double const hi[3] = {m_info.bounds_max[0], m_info.bounds_max[1], m_info.bounds_max[2]};
    double const lp[3] = {light->position[0], light->position[1], light->position[2]};
    double f[3] = {0.5 * (lo[0] + hi[0]) - lp[0], 0.5 * (lo[1] + hi[1]) - lp[1], 0.5 * (lo[2] + hi[2]) - lp[2]};
    double const fl = std::sqrt(f[0] * f[0] + f[1] * f[1] + f[2] * f[2]);
    double const diag = std::sqrt((hi[0] - lo[0]) * (hi[0] - lo[0]) + (hi[1] - lo[1]) * (hi[1] - lo[1])
                                  + (hi[2] - lo[2]) * (hi[2] - lo[2]));
    if (!(fl > 1e-6 * (diag + 1e-30)) || !std::isfinite(fl)) return;
    for (double& v : f) v /= fl;
    // any unit vector not parallel to f
    int const k = std::fabs(f[0]) <= std::fabs(f[1]) && std::fabs(f[0]) <= std::fabs(f[2]) ? 0
                : std::fabs(f[1]) <= std::fabs(f[2]) ? 1 : 2;
    double a[3] = {0, 0, 0};
    a[k] = 1.0;
    double r[3] = {a[1] * f[2] - a[2] * f[1], a[2] * f[0] - a[0] * f[2], a[0] * f[1] - a[1] * f[0]};
    double const rl = std::sqrt(r[0] * r[0] + r[1] * r[1] + r[2] * r[2]);
    for (double& v : r) v /= rl;
    double const u[3] = {f[1] * r[2] - f[2] * r[1], f[2] * r[0] - f[0] * r[2], f[0] * r[1] - f[1] * r[0]};
    // the 8 corners of the scene box must all lie well in front of the plane; their projections bound the grid
    double tmax_x = 0, tmax_y = 0;
    for (int c = 0; c < 8; c++)
    {
        double const p[3] = {((c & 1) ? hi[0] : lo[0]) - lp[0], ((c & 2) ? hi[1] : lo[1]) - lp[1],
                             ((c & 4) ? hi[2] : lo[2]) - lp[2]};
        double const cz = p[0] * f[0] + p[1] * f[1] + p[2] * f[2];
        if (!(cz > 1e-2 * (diag + fl))) return;
        tmax_x = std::max(tmax_x, std::fabs((p[0] * r[0] + p[1] * r[1] + p[2] * r[2]) / cz));
        tmax_y = std::max(tmax_y, std::fabs((p[0] * u[0] + p[1] * u[1] + p[2] * u[2]) / cz));
    }
    if (!(tmax_x > 0) || !(tmax_y > 0) || tmax_x > 50 || tmax_y > 50) return;
    double const half_w = tmax_x * 1.001 + 1e-6, half_h = tmax_y * 1.001 + 1e-6;
    // about 128 cells per triangle, split by the aspect of the projected scene: a cell's list then holds little
    // more than the triangles that really cover it.  Measured on the 100k-triangle frame (one B200, Grays/s |
    // one-time build ms): 16: 48.8 | 1.2, 32: 53.0 | 1.2, 64: 55.6, 128: 57.4 | 1.3, 256: 58.7 | 2.0, 512: 59.3 | 3.4
    // -- the walk's cost is the entries it visits that do NOT cover the point.  128 is where frame + build is least.
    static double const cells_per_tri = std::getenv("SRT_LIGHT_GRID_CELLS") ? std::atof(std::getenv("SRT_LIGHT_GRID_CELLS")) : 128.0;
    static double const max_side = std::getenv("SRT_LIGHT_GRID_MAX") ? std::atof(std::getenv("SRT_LIGHT_GRID_MAX")) : 4096.0;
    double const g = std::min(max_side, std::max(256.0, std::sqrt(cells_per_tri * static_cast<double>(m_view.n_tris))));
    uint32_t const gw = static_cast<uint32_t>(std::min(max_side, std::max(64.0, g * std::sqrt(half_w / half_h))));
    uint32_t const gh = static_cast<uint32_t>(std::min(max_side, std::max(64.0, g * std::sqrt(half_h / half_w))));
    uint32_t const tiles = gw * gh;
    // entries: a triangle is listed in every cell its projection touches, so the total grows with the cell count
    uint64_t const want_cap = std::max<uint64_t>(96ull * m_view.n_tris, 3ull * tiles);
    uint32_t const cap = std::max<uint32_t>(static_cast<uint32_t>(std::min<uint64_t>(want_cap, 1ull << 28)), 1u << 20);
    if (tiles > m_lgrid_tiles || cap > m_lgrid_cap || m_d_lgrid_bbox == nullptr)
    {
        ML_CUDA_CHECK(cudaStreamSynchronize(stream));
        if (m_grid_free) ML_CUDA_CHECK(cudaEventSynchronize(m_grid_free));
        cudaFree(m_d_lgrid_tiles); cudaFree(m_d_lgrid_items); cudaFree(m_d_lgrid_bbox); cudaFree(m_d_lgrid_misc);
        m_d_lgrid_tiles = m_d_lgrid_items = m_d_lgrid_bbox = m_d_lgrid_misc = nullptr;
        m_lgrid_tiles = m_lgrid_cap = 0;
        ML_CUDA_CHECK(cudaMalloc(&m_d_lgrid_tiles, (3 * static_cast<size_t>(tiles) + 8 + tiles / kGridScanCells + 8) * sizeof(uint32_t)));
        // tile counts are zero between builds (grid_scan_blocks_kernel re-zeroes them): no per-build memset
        ML_CUDA_CHECK(cudaMemsetAsync(m_d_lgrid_tiles, 0, (3 * static_cast<size_t>(tiles) + 8 + tiles / kGridScanCells + 8) * sizeof(uint32_t), stream));
        ML_CUDA_CHECK(cudaMalloc(&m_d_lgrid_items, static_cast<size_t>(cap) * sizeof(uint32_t)));
        ML_CUDA_CHECK(cudaMalloc(&m_d_lgrid_bbox, 4 * static_cast<size_t>(m_view.n_tris_padded) * sizeof(F4)));
        ML_CUDA_CHECK(cudaMalloc(&m_d_lgrid_misc, sizeof(GridFlags) + kGridMaxGlobal * sizeof(uint32_t)));
        m_lgrid_tiles = tiles;
        m_lgrid_cap = cap;
    }
    if (m_grid_free == nullptr) ML_CUDA_CHECK(cudaEventCreateWithFlags(&m_grid_free, cudaEventDisableTiming));
    ML_CUDA_CHECK(cudaStreamWaitEvent(stream, m_grid_free, 0));

    GridBuild b{};
    uint32_t* t = static_cast<uint32_t*>(m_d_lgrid_tiles);
    b.scene = m_view;
    for (int i = 0; i < 3; i++)
    {
        b.origin[i] = light->position[i];
        b.minv[i] = static_cast<float>(r[i]);          // orthonormal basis: the inverse is the transpose
        b.minv[3 + i] = static_cast<float>(u[i]);
        b.minv[6 + i] = static_cast<float>(f[i]);
    }
    b.half_w = static_cast<float>(half_w); b.half_h = static_cast<float>(half_h);
    b.pad = 4.0f * m_pad;
    b.width = gw; b.height = gh; b.tiles_w = gw; b.tiles_h = gh; b.cap = cap;
    b.tile = 1; b.lo_bias = 0.f;
    b.ty_begin = 0; b.ty_end = gh; b.ty_stride = 1; b.ty_phase = 0;
    // built once per light position: long walks are affordable.  A scene of at most kGridMaxGlobal triangles needs
    // no lists at all: every triangle goes to the global list and each shadow ray tests them all.
    b.max_tiles_per_tri = m_view.n_tris <= kGridMaxGlobal ? 0u : 1u << 16;
    // laid out by the CAPACITY of the allocation, so the count array never overlaps another array of an
    // earlier, differently sized build (its all-zero invariant replaces the per-build memset)
    size_t const cap_tiles = m_lgrid_tiles;
    b.tile_cnt = t; b.tile_off = t + cap_tiles; b.cursor = t + 2 * cap_tiles + 1;
    b.block_sums = t + 3 * cap_tiles + 4;
    b.items = static_cast<uint32_t*>(m_d_lgrid_items);
    b.bbox = static_cast<F4*>(m_d_lgrid_bbox);
    b.edges = b.bbox + m_view.n_tris_padded;
    b.half_extent = 0.f;
    b.flags = static_cast<GridFlags*>(m_d_lgrid_misc);
    b.global = reinterpret_cast<uint32_t*>(b.flags + 1);
    if (m_lgrid_built == nullptr)
    {
        ML_CUDA_CHECK(cudaEventCreate(&m_lgrid_ev0));     // timing enabled: srt_render_stats::light_grid_build_ms
        ML_CUDA_CHECK(cudaEventCreate(&m_lgrid_built));
    }
    ML_CUDA_CHECK(cudaEventRecord(m_lgrid_ev0, stream));
    BuildScreenGrid(b, stream);
    LaunchSortCellsByDepth(b, stream);     // shadow rays stop at the first entry beyond the shading point

    GridView& gv = m_lgrid.grid;
    gv.tile_off = b.tile_off; gv.items = b.items; gv.bbox = b.bbox; gv.edges = b.edges; gv.global = b.global;
    gv.flags = b.flags;
    gv.tiles_w = gw; gv.tiles_h = gh;
    for (int i = 0; i < 9; i++) m_lgrid.minv[i] = b.minv[i];
    m_lgrid.kx = 0.5f * static_cast<float>(gw) / b.half_w;
    m_lgrid.ky = 0.5f * static_cast<float>(gh) / b.half_h;
    m_lgrid.cx = 0.5f * static_cast<float>(gw);
    m_lgrid.cy = 0.5f * static_cast<float>(gh);
    static int const max_walk = std::getenv("SRT_LIGHT_WALK") ? std::atoi(std::getenv("SRT_LIGHT_WALK")) : 48;
    m_lgrid.max_walk = static_cast<uint32_t>(std::max(max_walk, 1));
    gv.enabled = 1;
    ML_CUDA_CHECK(cudaEventRecord(m_lgrid_built, stream));
    std::memcpy(m_lgrid_light, light->position, sizeof(m_lgrid_light));
    m_lgrid_valid = true;
    static int const debug = std::getenv("SRT_GRID_DEBUG") ? std::atoi(std::getenv("SRT_GRID_DEBUG")) : 0;
    if (debug)
    {
        GridFlags verdict{};
        ML_CUDA_CHECK(cudaStreamSynchronize(stream));
        ML_CUDA_CHECK(cudaMemcpy(&verdict, b.flags, sizeof(verdict), cudaMemcpyDeviceToHost));
        std::fprintf(stderr, "[srt grid] light %ux%u cells, entries %u (cap %u), global %u, ok %u\n", gw, gh, verdict.total,
                     cap, verdict.n_global, verdict.ok);
    }
}

// Primary-ray grid for this call's camera and frame, built on `stream` (scene.h, GridView).  Leaves
// m_grid.enabled = 0 when the render will not use it (kernel choice, SRT_GRID=0, singular camera basis).
void Scene::PrepareGrid(srt_camera const* camera, srt_render_params const* params, cudaStream_t stream)
{
    m_grid.enabled = 0;
    int const want = std::getenv("SRT_GRID") ? std::atoi(std::getenv("SRT_GRID")) : 1;   // read per call: tests flip it
    if (!want || params->accel != SRT_ACCEL_BVH || !(RenderUsesGrid(*params) & 1) || m_view.n_tris == 0) return;
    if (params->y0 % kGridTile != 0) return;     // tiles are the render kernels' 8-row blocks

    // inverse of the camera basis [right up forward] in double: P - o = cx*right + cy*up + cz*forward
    double const a[3][3] = {{camera->right[0], camera->up[0], camera->forward[0]},
                            {camera->right[1], camera->up[1], camera->forward[1]},
                            {camera->right[2], camera->up[2], camera->forward[2]}};
    double const det = a[0][0] * (a[1][1] * a[2][2] - a[1][2] * a[2][1])
                     - a[0][1] * (a[1][0] * a[2][2] - a[1][2] * a[2][0])
                     + a[0][2] * (a[1][0] * a[2][1] - a[1][1] * a[2][0]);
    double scale = 0;
    for (auto const& r : a) for (double v : r) scale = std::max(scale, std::fabs(v));
    if (!(std::fabs(det) > 1e-6 * scale * scale * scale) || !(camera->half_w > 0.f) || !(camera->half_h > 0.f)) return;
    GridBuild b{};
    double inv[3][3];
    inv[0][0] = (a[1][1] * a[2][2] - a[1][2] * a[2][1]) / det;
    inv[0][1] = (a[0][2] * a[2][1] - a[0][1] * a[2][2]) / det;
    inv[0][2] = (a[0][1] * a[1][2] - a[0][2] * a[1][1]) / det;
    inv[1][0] = (a[1][2] * a[2][0] - a[1][0] * a[2][2]) / det;
    inv[1][1] = (a[0][0] * a[2][2] - a[0][2] * a[2][0]) / det;
    inv[1][2] = (a[0][2] * a[1][0] - a[0][0] * a[1][2]) / det;
    inv[2][0] = (a[1][0] * a[2][1] - a[1][1] * a[2][0]) / det;
    inv[2][1] = (a[0][1] * a[2][0] - a[0][0] * a[2][1]) / det;
    inv[2][2] = (a[0][0] * a[1][1] - a[0][1] * a[1][0]) / det;
    for (int r = 0; r < 3; r++)
        for (int c = 0; c < 3; c++)
        {
            if (!std::isfinite(inv[r][c])) return;
            b.minv[3 * r + c] = static_cast<float>(inv[r][c]);
        }

    uint32_t const tw = (params->width + kGridTile - 1) / kGridTile, th = (params->height + kGridTile - 1) / kGridTile;
    uint32_t const tiles = tw * th;
    uint32_t const cap = std::max<uint32_t>(16u * m_view.n_tris, 1u << 20);
    if (tiles > m_grid_tiles || cap > m_grid_cap || m_d_grid_bbox == nullptr)
    {
        ML_CUDA_CHECK(cudaStreamSynchronize(stream));
        if (m_grid_free) ML_CUDA_CHECK(cudaEventSynchronize(m_grid_free));
        FreeGrid();
        ML_CUDA_CHECK(cudaMalloc(&m_d_grid_tiles, (3 * static_cast<size_t>(tiles) + 8 + tiles / kGridScanCells + 8) * sizeof(uint32_t)));
        ML_CUDA_CHECK(cudaMemsetAsync(m_d_grid_tiles, 0, (3 * static_cast<size_t>(tiles) + 8 + tiles / kGridScanCells + 8) * sizeof(uint32_t), stream));
        ML_CUDA_CHECK(cudaMalloc(&m_d_grid_items, static_cast<size_t>(cap) * sizeof(uint32_t)));
        ML_CUDA_CHECK(cudaMalloc(&m_d_grid_bbox, 4 * static_cast<size_t>(m_view.n_tris_padded) * sizeof(F4)));
        ML_CUDA_CHECK(cudaMalloc(&m_d_grid_misc, sizeof(GridFlags) + kGridMaxGlobal * sizeof(uint32_t)));
        // per-tile times of the previous frame (zero = no history) | up to 8 items per tile and class
        ML_CUDA_CHECK(cudaMalloc(&m_d_grid_order, (static_cast<size_t>(kGridBuckets) * 8u + 1u) * tiles * sizeof(uint32_t)));
        ML_CUDA_CHECK(cudaMemsetAsync(m_d_grid_order, 0, static_cast<size_t>(tiles) * sizeof(uint32_t), stream));
        m_grid_tiles = tiles;
        m_grid_cap = cap;
    }
    if (m_grid_free == nullptr) ML_CUDA_CHECK(cudaEventCreateWithFlags(&m_grid_free, cudaEventDisableTiming));
    // a render that still reads the previous grid (possibly on another stream) must finish first
    ML_CUDA_CHECK(cudaStreamWaitEvent(stream, m_grid_free, 0));

    uint32_t* t = static_cast<uint32_t*>(m_d_grid_tiles);
    b.scene = m_view;
    for (int k = 0; k < 3; k++) b.origin[k] = camera->origin[k];
    b.half_w = camera->half_w; b.half_h = camera->half_h;
    b.pad = 4.0f * m_pad;
    b.width = params->width; b.height = params->height; b.tiles_w = tw; b.tiles_h = th; b.cap = cap;
    b.tile = kGridTile; b.lo_bias = 1.f;
    // only the tile rows this call renders (a rank of an interleaved multi-GPU frame owns every N-th)
    b.ty_begin = params->y0 / kGridTile;
    b.ty_end = (params->y1 + kGridTile - 1) / kGridTile;
    b.ty_stride = params->row_block_stride > 1 ? params->row_block_stride : 1;
    b.ty_phase = params->row_block_phase;
    b.max_tiles_per_tri = m_view.n_tris <= kGridMaxGlobal ? 0u : 1024u;   // tiny scenes: all global, no binning
    // laid out by the CAPACITY of the allocation, so the count array never overlaps another array of an
    // earlier, differently sized build (its all-zero invariant replaces the per-build memset)
    size_t const cap_tiles = m_grid_tiles;
    b.tile_cnt = t; b.tile_off = t + cap_tiles; b.cursor = t + 2 * cap_tiles + 1;
    b.block_sums = t + 3 * cap_tiles + 4;
    b.items = static_cast<uint32_t*>(m_d_grid_items);
    b.bbox = static_cast<F4*>(m_d_grid_bbox);
    b.edges = b.bbox + m_view.n_tris_padded;
    b.half_extent = 0.5f;
    b.flags = reinterpret_cast<GridFlags*>(reinterpret_cast<uint32_t*>(m_d_counters) + kScratchGridFlagsWord);
    b.flags_prezeroed = 1;     // cleared by the render call's scratch memset
    b.global = reinterpret_cast<uint32_t*>(static_cast<GridFlags*>(m_d_grid_misc) + 1);
    b.tile_time = static_cast<uint32_t*>(m_d_grid_order);
    b.order = b.tile_time + m_grid_tiles;
    b.bucket_cnt = reinterpret_cast<uint32_t*>(m_d_counters) + kScratchBucketWord;
    b.order_stride = 8u * m_grid_tiles;
    BuildScreenGrid(b, stream);

    m_grid.tile_off = b.tile_off; m_grid.items = b.items; m_grid.bbox = b.bbox; m_grid.edges = b.edges;
    m_grid.global = b.global;
    m_grid.flags = b.flags; m_grid.tiles_w = tw; m_grid.tiles_h = th;
    m_grid.order = b.order; m_grid.bucket_cnt = b.bucket_cnt; m_grid.order_stride = b.order_stride;
    m_grid.tile_time = b.tile_time;
    m_grid.enabled = 1;
    static int const debug = std::getenv("SRT_GRID_DEBUG") ? std::atoi(std::getenv("SRT_GRID_DEBUG")) : 0;
    if (debug)
    {
        GridFlags f{};
        ML_CUDA_CHECK(cudaStreamSynchronize(stream));
        ML_CUDA_CHECK(cudaMemcpy(&f, b.flags, sizeof(f), cudaMemcpyDeviceToHost));
        std::fprintf(stderr, "[srt grid] screen %ux%u tiles, entries %u (cap %u), global %u, ok %u\n", tw, th, f.total, cap,
                     f.n_global, f.ok);
    }
}

void Scene::FreeStaging()
{
    cudaFree(m_d_rgb); cudaFree(m_d_id); cudaFree(m_d_t);
    cudaFreeHost(m_h_rgb); cudaFreeHost(m_h_id); cudaFreeHost(m_h_t);
    m_d_rgb = m_d_id = m_d_t = m_h_rgb = m_h_id = m_h_t = nullptr;
    m_stage_pixels = 0;
}

Scene::~Scene()
{
    cudaSetDevice(m_device->device);
    cudaStreamSynchronize(m_device->stream);
    FreeStaging();
    FreeGrid();
    if (m_grid_free) cudaEventDestroy(m_grid_free);
    if (m_lgrid_built) cudaEventDestroy(m_lgrid_built);
    if (m_lgrid_ev0) cudaEventDestroy(m_lgrid_ev0);
    if (m_ev_grid_done) cudaEventDestroy(m_ev_grid_done);
    if (m_ev_fork) cudaEventDestroy(m_ev_fork);
    if (m_ev_join) cudaEventDestroy(m_ev_join);
    for (cudaEvent_t e : m_chunk_events) cudaEventDestroy(e);
    cudaFree(m_d_tris); cudaFree(m_d_albedo); cudaFree(m_d_nodes); cudaFree(m_d_counters);
}

ml_status Scene::GetInfo(srt_scene_info* info) const
{
    if (info == nullptr) return ML_FAIL;
    *info = m_info;
    return ML_OK;
}

size_t Scene::GetBvh(float* nodes, size_t max_nodes, int32_t* order, size_t max_order) const
{
    if (nodes)
        std::memcpy(nodes, m_nodes_host.data(), std::min(max_nodes, m_nodes_host.size()) * sizeof(BvhNode));
    if (order)
        std::memcpy(order, m_order_host.data(), std::min(max_order, m_order_host.size()) * sizeof(int32_t));
    return m_nodes_host.size();
}

bool Scene::Validate(srt_camera const* camera, srt_light const* light, srt_render_params const* p)
{
    m_error.clear();
    std::ostringstream os;
    if (camera == nullptr) os << "Bad camera argument";
    else if (light == nullptr) os << "Bad light argument";
    else if (p == nullptr) os << "Bad render parameters argument";
    else if (p->width == 0 || p->height == 0) os << "Unspecified frame dimension";
    else if (static_cast<uint64_t>(p->width) * p->height >= (uint64_t(1) << 31)) os << "Frame too large";
    else if (p->y0 > p->y1 || p->y1 > p->height)
        os << "Bad row band [" << p->y0 << ", " << p->y1 << ") for height " << p->height;
    else if (p->buf_y0 > p->y0) os << "Bad buffer row origin " << p->buf_y0 << " for band start " << p->y0;
    else if (p->row_block_stride > 1 && (p->y0 % 8 != 0 || p->row_block_phase >= p->row_block_stride))
        os << "Bad interleaved band: y0 " << p->y0 << " must be a multiple of 8 and phase "
           << p->row_block_phase << " below stride " << p->row_block_stride;
    else if (p->spp == 0) os << "Unspecified sample count";
    else if (p->shade > SRT_SHADE_SHADOW) os << "Unsupported shade mode: " << p->shade;
    else if (p->accel > SRT_ACCEL_BVH) os << "Unsupported acceleration mode: " << p->accel;
    else if (p->accel == SRT_ACCEL_BVH && !HasBvh()) os << "Scene was created without a BVH";
    m_error = os.str();
    return m_error.empty();
}

ml_status Scene::RenderDevice(srt_camera const* camera, srt_light const* light,
                              srt_render_params const* params, void* d_rgb, void* d_hit_id,
                              void* d_hit_t, cudaStream_t stream, bool timed, SignalSpec const* signal)
{
    if (!Validate(camera, light, params)) return ML_FAIL;
    try
    {
        m_device->MakeCurrent();
        if (stream == nullptr) stream = m_device->stream;
        if (m_ev_grid_done == nullptr)
        {
            ML_CUDA_CHECK(cudaEventCreate(&m_ev_grid_done));
            ML_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_fork, cudaEventDisableTiming));
            ML_CUDA_CHECK(cudaEventCreateWithFlags(&m_ev_join, cudaEventDisableTiming));
        }
        ML_CUDA_CHECK(cudaMemsetAsync(m_d_counters, 0, kScratchBytes, stream));   // counters, tickets, screen-grid flags
        if (timed) ML_CUDA_CHECK(cudaEventRecord(m_device->ev_start, stream));
        PrepareGrid(camera, params, stream);       // inside the timed region: it is per-frame work
        PrepareLightGrid(light, params, stream);
        if (timed) ML_CUDA_CHECK(cudaEventRecord(m_ev_grid_done, stream));
        uint32_t chunk_rows = params->y1 - params->y0, n_chunks = 1;
        if (signal && signal->n_chunks > 1) ChunkLayout(*params, signal->n_chunks, chunk_rows, n_chunks);
        if (n_chunks <= 1)
        {
            EnqueueBand(camera, light, params, d_rgb, d_hit_id, d_hit_t, stream,
                        signal ? signal->d_flags : nullptr, signal ? signal->value : 0u, 0, true);
        }
        else
        {
            // Chunk launches alternate between the caller's stream and a helper stream, so chunk k+1's blocks
            // fill the SMs that chunk k's last wave leaves idle; flag k is published by chunk k's last block.
            cudaStream_t const s1 = m_device->copy_in;
            ML_CUDA_CHECK(cudaEventRecord(m_ev_fork, stream));
            ML_CUDA_CHECK(cudaStreamWaitEvent(s1, m_ev_fork, 0));
            for (uint32_t k = 0; k < n_chunks; k++)
            {
                srt_render_params c = *params;
                c.y0 = params->y0 + k * chunk_rows;
                c.y1 = std::min(params->y1, c.y0 + chunk_rows);
                EnqueueBand(camera, light, &c, d_rgb, d_hit_id, d_hit_t, (k & 1u) ? s1 : stream,
                            signal->d_flags + static_cast<size_t>(k) * signal->chunk_stride, signal->value, k & 1u, false);
            }
            ML_CUDA_CHECK(cudaEventRecord(m_ev_join, s1));
            ML_CUDA_CHECK(cudaStreamWaitEvent(stream, m_ev_join, 0));
        }
        if (m_grid.enabled || m_lgrid.grid.enabled) ML_CUDA_CHECK(cudaEventRecord(m_grid_free, stream));
        if (timed) ML_CUDA_CHECK(cudaEventRecord(m_device->ev_stop, stream));
        ML_CUDA_CHECK(cudaGetLastError());
        m_timed_pending = timed;
        m_timed_grid = timed;
        m_timed_stream = stream;
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_error = e.what();
        return ML_FAIL;
    }
}

// One kernel launch for the rows params->[y0, y1) (already validated); counters accumulate.  With d_signal
// the launch's last block publishes `value` there (RenderLaunch::d_signal); a band in which this call owns
// no row block still publishes it, from a one-thread kernel.
void Scene::EnqueueBand(srt_camera const* camera, srt_light const* light, srt_render_params const* params,
                        void* d_rgb, void* d_hit_id, void* d_hit_t, cudaStream_t stream,
                        uint32_t* d_signal, uint32_t signal_value, uint32_t ticket, bool whole_band)
{
    RenderLaunch launch{};
    launch.scene = m_view;
    launch.camera = *camera;
    launch.light = *light;
    launch.params = *params;
    launch.inv2w = 2.0f / static_cast<float>(params->width);
    launch.inv2h = 2.0f / static_cast<float>(params->height);
    launch.inv_spp = 1.0f / static_cast<float>(params->spp);
    launch.d_rgb = static_cast<float*>(d_rgb);
    launch.d_hit_id = static_cast<int32_t*>(d_hit_id);
    launch.d_hit_t = static_cast<float*>(d_hit_t);
    launch.d_counters = m_d_counters;
    launch.d_signal = d_signal;
    launch.signal_value = signal_value;
    launch.d_ticket = reinterpret_cast<uint32_t*>(m_d_counters) + kScratchTicketWord + (ticket & 1u);
    static int const trace = std::getenv("SRT_BLOCK_TRACE") ? std::atoi(std::getenv("SRT_BLOCK_TRACE")) : 0;
    launch.d_trace = trace ? m_d_counters + kScratchTraceU64 : nullptr;
    RowBlocks(*params, launch.row_origin, launch.row_block_step, launch.n_row_blocks);
    static int const hints = std::getenv("SRT_BVH_HINTS") ? std::atoi(std::getenv("SRT_BVH_HINTS")) : 1;
    launch.hints = static_cast<uint32_t>(hints);
    launch.grid = m_grid;
    if (launch.row_origin % kGridTile != 0) launch.grid.enabled = 0;
    launch.lgrid = m_lgrid;
    // heaviest-first work list: only a launch over exactly the band the grid (and its list) was built for
    // Measured (config 3): the work list costs ~3 % of a single GPU's frame (4.91 vs 4.75 ms: item fetch, two
    // barriers per item) and wins as soon as the frame is sharded (N = 2: 2.57 vs 2.63 ms; N = 8: long tiles no
    // longer bound the rank).  SRT_BLOCK_ORDER: 0 never, 1 sharded frames (default), 2 always.
    static int const want_order = std::getenv("SRT_BLOCK_ORDER") ? std::atoi(std::getenv("SRT_BLOCK_ORDER")) : 1;
    bool const sharded = params->row_block_stride > 1;
    launch.ordered = (whole_band && (want_order == 2 || (want_order == 1 && sharded)) && launch.grid.enabled
                      && launch.grid.order != nullptr) ? 1u : 0u;
    launch.d_work = reinterpret_cast<uint32_t*>(m_d_counters) + kScratchWorkWord;
    if (launch.n_row_blocks > 0)
        LaunchRender(launch, m_device->prop.multiProcessorCount, stream);
    else if (d_signal)
        LaunchWriteFlag(d_signal, signal_value, stream);
}

// srtRenderDeviceNoAlloc.  A separate path on purpose (RenderDevice is untouched): nothing here allocates, creates
// events or touches the scene's own grid buffers; the per-frame scratch is the static array of render_kernels.cu.
ml_status Scene::RenderDeviceNoAlloc(srt_camera const* camera, srt_light const* light, srt_render_params const* params,
                                     void* d_rgb, void* d_hit_id, void* d_hit_t, cudaStream_t stream)
{
    if (!Validate(camera, light, params)) return ML_FAIL;
    try
    {
        uint32_t const tw = (params->width + kGridTile - 1) / kGridTile, th = (params->height + kGridTile - 1) / kGridTile;
        if (m_view.n_tris > SRT_NOALLOC_MAX_TRIANGLES || m_view.n_tris_padded > SRT_NOALLOC_MAX_TRIANGLES)
            throw std::runtime_error("srtRenderDeviceNoAlloc: the scene has " + std::to_string(m_view.n_tris)
                                     + " triangles; the static scratch holds at most "
                                     + std::to_string(SRT_NOALLOC_MAX_TRIANGLES) + " (use srtRenderDevice)");
        if (static_cast<uint64_t>(tw) * th > SRT_NOALLOC_MAX_TILES)
            throw std::runtime_error("srtRenderDeviceNoAlloc: a " + std::to_string(params->width) + " x "
                                     + std::to_string(params->height) + " frame has " + std::to_string(uint64_t(tw) * th)
                                     + " 8x8 tiles; the static scratch holds at most "
                                     + std::to_string(SRT_NOALLOC_MAX_TILES) + " (use srtRenderDevice)");
        m_device->MakeCurrent();
        if (stream == nullptr) stream = m_device->stream;

        RenderLaunch launch{};
        launch.scene = m_view;
        launch.camera = *camera;
        launch.light = *light;
        launch.params = *params;
        launch.params.collect_stats = 0;
        launch.inv2w = 2.0f / static_cast<float>(params->width);
        launch.inv2h = 2.0f / static_cast<float>(params->height);
        launch.inv_spp = 1.0f / static_cast<float>(params->spp);
        launch.d_rgb = static_cast<float*>(d_rgb);
        launch.d_hit_id = static_cast<int32_t*>(d_hit_id);
        launch.d_hit_t = static_cast<float*>(d_hit_t);
        RowBlocks(*params, launch.row_origin, launch.row_block_step, launch.n_row_blocks);
        launch.hints = 1;

        // screen grid for this camera, as PrepareGrid decides and parameterises it (pointers: RenderNoAllocScratch)
        GridBuild b{};
        bool use_grid = params->accel == SRT_ACCEL_BVH && (RenderUsesGrid(*params) & 1) && m_view.n_tris > 0
                        && params->y0 % kGridTile == 0 && launch.row_origin % kGridTile == 0;
        if (use_grid)
        {
            double const a[3][3] = {{camera->right[0], camera->up[0], camera->forward[0]},
                                    {camera->right[1], camera->up[1], camera->forward[1]},
                                    {camera->right[2], camera->up[2], camera->forward[2]}};
            double const c00 = a[1][1] * a[2][2] - a[1][2] * a[2][1], c01 = a[1][0] * a[2][2] - a[1][2] * a[2][0],
                         c02 = a[1][0] * a[2][1] - a[1][1] * a[2][0];
            double const det = a[0][0] * c00 - a[0][1] * c01 + a[0][2] * c02;
            double scale = 0;
            for (auto const& r : a) for (double v : r) scale = std::max(scale, std::fabs(v));
            use_grid = std::fabs(det) > 1e-6 * scale * scale * scale && camera->half_w > 0.f && camera->half_h > 0.f;
            double const inv[9] = {c00 / det, (a[0][2] * a[2][1] - a[0][1] * a[2][2]) / det,
                                   (a[0][1] * a[1][2] - a[0][2] * a[1][1]) / det,
                                   -c01 / det, (a[0][0] * a[2][2] - a[0][2] * a[2][0]) / det,
                                   (a[0][2] * a[1][0] - a[0][0] * a[1][2]) / det,
                                   c02 / det, (a[0][1] * a[2][0] - a[0][0] * a[2][1]) / det,
                                   (a[0][0] * a[1][1] - a[0][1] * a[1][0]) / det};
            for (int k = 0; k < 9 && use_grid; k++)
            {
                use_grid = std::isfinite(inv[k]);
                b.minv[k] = static_cast<float>(inv[k]);
            }
        }
        if (use_grid)
        {
            b.scene = m_view;
            for (int k = 0; k < 3; k++) b.origin[k] = camera->origin[k];
            b.half_w = camera->half_w; b.half_h = camera->half_h;
            b.pad = 4.0f * m_pad;
            b.width = params->width; b.height = params->height; b.tiles_w = tw; b.tiles_h = th;
            b.tile = kGridTile; b.lo_bias = 1.f;
            b.ty_begin = params->y0 / kGridTile;
            b.ty_end = (params->y1 + kGridTile - 1) / kGridTile;
            b.ty_stride = params->row_block_stride > 1 ? params->row_block_stride : 1;
            b.ty_phase = params->row_block_phase;
            b.max_tiles_per_tri = m_view.n_tris <= kGridMaxGlobal ? 0u : 1024u;
            b.half_extent = 0.5f;
        }
        RenderNoAllocScratch(m_device->device, launch, b, use_grid, m_device->prop.multiProcessorCount, stream);
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_error = e.what();
        return ML_FAIL;
    }
}

ml_status Scene::GetStats(srt_render_stats* stats)
{
    if (stats == nullptr) return ML_FAIL;
    try
    {
        m_device->MakeCurrent();
        cudaStream_t s = m_timed_stream ? m_timed_stream : m_device->stream;
        unsigned long long c[4] = {};
        ML_CUDA_CHECK(cudaMemcpyAsync(c, m_d_counters, sizeof(c), cudaMemcpyDeviceToHost, s));
        ML_CUDA_CHECK(cudaStreamSynchronize(s));
        stats->primary_rays = c[0];
        stats->shadow_rays = c[1];
        stats->box_tests = c[2];
        stats->tri_tests = c[3];
        stats->kernel_ms = 0.f;
        stats->grid_build_ms = 0.f;
        stats->light_grid_build_ms = 0.f;
        if (m_timed_pending)
            ML_CUDA_CHECK(cudaEventElapsedTime(&stats->kernel_ms, m_device->ev_start, m_device->ev_stop));
        if (m_timed_pending && m_timed_grid)
            ML_CUDA_CHECK(cudaEventElapsedTime(&stats->grid_build_ms, m_device->ev_start, m_ev_grid_done));
        if (m_lgrid_valid && m_lgrid_ev0 != nullptr && cudaEventQuery(m_lgrid_built) == cudaSuccess)
            ML_CUDA_CHECK(cudaEventElapsedTime(&stats->light_grid_build_ms, m_lgrid_ev0, m_lgrid_built));
        if (std::getenv("SRT_BLOCK_TRACE") && std::atoi(std::getenv("SRT_BLOCK_TRACE")))
        {
            // diagnostic: when did the launch run out of fresh blocks, and how long did it drain
            unsigned long long t[8] = {};
            ML_CUDA_CHECK(cudaMemcpy(t, m_d_counters + kScratchTraceU64, sizeof(t), cudaMemcpyDeviceToHost));
            unsigned long long const first = ~t[0];
            if (t[2] > first)
                std::fprintf(stderr, "[srt trace] dev %d events: call %.1f us, grid build %.1f us | items: span %.1f us, last start at %.1f us (drain %.1f us), "
                                     "sum of block durations %.1f us, longest block %.1f us\n",
                             m_device->device, stats->kernel_ms * 1e3, stats->grid_build_ms * 1e3,
                             (t[2] - first) * 1e-3, (t[1] - first) * 1e-3, (t[2] - t[1]) * 1e-3, t[3] * 1e-3, t[4] * 1e-3);
            {
                unsigned long long e[2] = {};
                ML_CUDA_CHECK(cudaMemcpy(e, m_d_counters + kScratchTraceU64 + 12, sizeof(e), cudaMemcpyDeviceToHost));
                if (e[0] && e[1])
                    std::fprintf(stderr, "[srt trace] dev %d kernel: first block entry -> first item %.1f us, last item end -> "
                                         "last block exit %.1f us\n", m_device->device, (static_cast<double>(first) - static_cast<double>(~e[0])) * 1e-3,
                                 (static_cast<double>(e[1]) - static_cast<double>(t[2])) * 1e-3);
            }
            if (t[5] && m_grid.tiles_w)
            {
                uint32_t buckets[9] = {};
                ML_CUDA_CHECK(cudaMemcpy(buckets, reinterpret_cast<uint32_t*>(m_d_counters) + kScratchBucketWord,
                                         8 * sizeof(uint32_t), cudaMemcpyDeviceToHost));
                ML_CUDA_CHECK(cudaMemcpy(buckets + 8, reinterpret_cast<uint32_t*>(m_d_counters) + kScratchWorkWord,
                                         sizeof(uint32_t), cudaMemcpyDeviceToHost));
                unsigned const tile = static_cast<unsigned>(t[5] & 0xffffffu);
                std::fprintf(stderr, "[srt trace] longest item: %.1f us, tile (%u, %u), list length %u; items per class "
                                     "%u / %u / %u / %u / %u / %u / %u / %u, cursor %u\n", (t[5] >> 34) * 1e-3,
                             tile % m_grid.tiles_w, tile / m_grid.tiles_w, static_cast<unsigned>((t[5] >> 24) & 1023u),
                             buckets[0], buckets[1], buckets[2], buckets[3], buckets[4], buckets[5], buckets[6], buckets[7],
                             buckets[8]);
                if (t[6] || t[7])
                    std::fprintf(stderr, "[srt trace]   its warp 0: box tests max lane %llu, all lanes %llu; triangle tests max "
                                         "lane %llu, all lanes %llu; lane 0 cast %llu shadow rays\n", t[6] >> 32,
                                 t[6] & 0xffffffffull, (t[7] >> 20) & 0xfffffull, t[7] & 0xfffffull, t[7] >> 40);
            }
        }

        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_error = e.what();
        return ML_FAIL;
    }
}

void Scene::EnsureStaging(size_t pixels)
{
    if (pixels <= m_stage_pixels) return;
    FreeStaging();
    ML_CUDA_CHECK(cudaMalloc(&m_d_rgb, pixels * 12));
    ML_CUDA_CHECK(cudaMalloc(&m_d_id, pixels * 4));
    ML_CUDA_CHECK(cudaMalloc(&m_d_t, pixels * 4));
    ML_CUDA_CHECK(cudaMallocHost(&m_h_rgb, pixels * 12));
    ML_CUDA_CHECK(cudaMallocHost(&m_h_id, pixels * 4));
    ML_CUDA_CHECK(cudaMallocHost(&m_h_t, pixels * 4));
    m_stage_pixels = pixels;
}

ml_status Scene::RenderHost(srt_camera const* camera, srt_light const* light,
                            srt_render_params const* params, float* rgb, int32_t* hit_id,
                            float* hit_t, srt_render_stats* stats)
{
    if (!Validate(camera, light, params)) return ML_FAIL;
    try
    {
        m_device->MakeCurrent();
        cudaStream_t const s0 = m_device->stream, s1 = m_device->copy_in, sc = m_device->copy_out;
        size_t const W = params->width;
        size_t const rows = params->y1 - params->y0;
        uint32_t const stride = params->row_block_stride > 1 ? params->row_block_stride : 1u;
        // staging holds whole 8-row blocks so a clipped last block can be copied as a unit
        EnsureStaging(std::max<size_t>((rows + 7) / 8 * 8 * W, 1));

        // The band is rendered into band-local device buffers (buf_y0 = y0) as up to 8 chunks of whole
        // (8 * stride)-row groups.  Chunk kernels alternate between two streams (chunk k+1's blocks fill
        // the SMs chunk k's last wave leaves idle); chunk k's D2H (copy stream, pinned staging) and its
        // copy into the caller's pageable frame run while later chunks render, so only the last chunk's
        // transfer is exposed.  With interleaved sharding only the owned 8-row blocks cross PCIe.
        srt_render_params band = *params;
        band.buf_y0 = params->y0;
        // Caller buffers that are page-locked (cudaHostAlloc / cudaHostRegister, as a whole) take the chunks
        // straight from the device: no staging copy, no host-side scatter.
        auto page_locked = [](void const* ptr) {
            if (ptr == nullptr) return true;
            cudaPointerAttributes at{};
            if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) { cudaGetLastError(); return false; }
            return at.type == cudaMemoryTypeHost;
        };
        bool const direct = (rgb || hit_id || hit_t) && page_locked(rgb) && page_locked(hit_id) && page_locked(hit_t);
        // band-local pixel p of the device buffers is pixel `to_caller + p` of the caller's frame
        size_t const to_caller = static_cast<size_t>(params->y0 - params->buf_y0) * W;
        void* const h_rgb = direct ? static_cast<void*>(rgb ? rgb + to_caller * 3 : nullptr) : m_h_rgb;
        void* const h_id = direct ? static_cast<void*>(hit_id ? hit_id + to_caller : nullptr) : m_h_id;
        void* const h_t = direct ? static_cast<void*>(hit_t ? hit_t + to_caller : nullptr) : m_h_t;
        size_t const group = 8u * stride;
        // chunks are sized by the rows this launch OWNS: each one costs a launch, two events and a copy
        // setup (~20 us), which a rank that owns 135 rows of a 1080p frame cannot hide behind 8 of them
        size_t const owned = rows / stride;
        static size_t const max_chunks = std::getenv("SRT_HOST_CHUNKS") ? std::max(1, std::atoi(std::getenv("SRT_HOST_CHUNKS"))) : 8;
        size_t n_chunks = owned >= 64 ? std::min<size_t>(max_chunks, std::max<size_t>(1, owned / 48)) : 1;
        size_t const chunk_rows = ((rows + n_chunks - 1) / n_chunks + group - 1) / group * group;
        n_chunks = (rows + chunk_rows - 1) / chunk_rows;
        while (m_chunk_events.size() < 2 * n_chunks + 2)
        {
            cudaEvent_t e = nullptr;
            ML_CUDA_CHECK(cudaEventCreateWithFlags(&e, cudaEventDisableTiming));
            m_chunk_events.push_back(e);
        }
        cudaEvent_t const ev_begin = m_chunk_events[2 * n_chunks], ev_s1_done = m_chunk_events[2 * n_chunks + 1];
        ML_CUDA_CHECK(cudaMemsetAsync(m_d_counters, 0, kScratchBytes, s0));
        ML_CUDA_CHECK(cudaEventRecord(m_device->ev_start, s0));
        PrepareGrid(camera, params, s0);
        PrepareLightGrid(light, params, s0);
        ML_CUDA_CHECK(cudaEventRecord(ev_begin, s0));
        ML_CUDA_CHECK(cudaStreamWaitEvent(s1, ev_begin, 0));
        for (size_t k = 0; k < n_chunks; k++)
        {
            cudaStream_t const s = (k & 1) ? s1 : s0;
            size_t const r0 = k * chunk_rows, r1 = std::min(rows, r0 + chunk_rows);
            srt_render_params c = band;
            c.y0 = params->y0 + static_cast<uint32_t>(r0);
            c.y1 = params->y0 + static_cast<uint32_t>(r1);
            EnqueueBand(camera, light, &c, m_d_rgb, m_d_id, m_d_t, s, nullptr, 0u, 0u, n_chunks == 1);
            ML_CUDA_CHECK(cudaEventRecord(m_chunk_events[2 * k], s));
            ML_CUDA_CHECK(cudaStreamWaitEvent(sc, m_chunk_events[2 * k], 0));
            if (stride == 1)
            {
                size_t const p0 = r0 * W, np = (r1 - r0) * W;
                auto copy = [&](void* h, void* d, size_t item) {
                    ML_CUDA_CHECK(cudaMemcpyAsync(static_cast<char*>(h) + p0 * item, static_cast<char*>(d) + p0 * item,
                                                  np * item, cudaMemcpyDeviceToHost, sc));
                };
                if (rgb) copy(h_rgb, m_d_rgb, 12);
                if (hit_id) copy(h_id, m_d_id, 4);
                if (hit_t) copy(h_t, m_d_t, 4);
            }
            else
            {
                // the chunk's owned 8-row blocks: one strided 2-D copy per buffer
                uint32_t origin, step, count;
                RowBlocks(c, origin, step, count);
                size_t const first = static_cast<size_t>(origin - params->y0) * W; // pixels
                size_t const blk = 8 * W;                                           // pixels per block
                size_t const pitch = static_cast<size_t>(step) * W;                 // pixels between blocks
                // the staging buffers hold whole blocks; the caller's frame may end inside the last one
                size_t const last_y = static_cast<size_t>(origin) + static_cast<size_t>(count ? count - 1 : 0) * step;
                bool const clipped = direct && count > 0 && last_y + 8 > c.y1;
                size_t const whole = clipped ? count - 1 : count;
                auto copy = [&](void* h, void* d, size_t item) {
                    if (whole > 0)
                        ML_CUDA_CHECK(cudaMemcpy2DAsync(static_cast<char*>(h) + first * item, pitch * item,
                                                        static_cast<char*>(d) + first * item, pitch * item,
                                                        blk * item, whole, cudaMemcpyDeviceToHost, sc));
                    if (clipped)
                    {
                        size_t const off = (last_y - params->y0) * W, n = (c.y1 - last_y) * W;
                        ML_CUDA_CHECK(cudaMemcpyAsync(static_cast<char*>(h) + off * item, static_cast<char*>(d) + off * item,
                                                      n * item, cudaMemcpyDeviceToHost, sc));
                    }
                };
                if (count > 0)
                {
                    if (rgb) copy(h_rgb, m_d_rgb, 12);
                    if (hit_id) copy(h_id, m_d_id, 4);
                    if (hit_t) copy(h_t, m_d_t, 4);
                }
            }
            ML_CUDA_CHECK(cudaEventRecord(m_chunk_events[2 * k + 1], sc));
        }
        ML_CUDA_CHECK(cudaEventRecord(ev_s1_done, s1));
        ML_CUDA_CHECK(cudaStreamWaitEvent(s0, ev_s1_done, 0));
        if (m_grid.enabled || m_lgrid.grid.enabled) ML_CUDA_CHECK(cudaEventRecord(m_grid_free, s0));
        ML_CUDA_CHECK(cudaEventRecord(m_device->ev_stop, s0));
        ML_CUDA_CHECK(cudaGetLastError());
        m_timed_pending = true;
        m_timed_grid = false;
        m_timed_stream = s0;

        // host side: as each chunk lands in the pinned staging, move its (owned) rows into the caller's
        // frame at their full-frame offsets (relative to the caller's buf_y0)
        for (size_t k = 0; k < n_chunks && !direct; k++)
        {
            size_t const r0 = k * chunk_rows, r1 = std::min(rows, r0 + chunk_rows);
            ML_CUDA_CHECK(cudaEventSynchronize(m_chunk_events[2 * k + 1]));
            for (size_t r = r0; r < r1;)
            {
                uint32_t const y = params->y0 + static_cast<uint32_t>(r);
                size_t run = r1 - r; // rows to copy in one go
                if (stride > 1)
                {
                    size_t const to_block_end = 8u - (y & 7u);
                    if ((y / 8u) % stride != params->row_block_phase) { r += to_block_end; continue; }
                    run = std::min(run, to_block_end);
                }
                size_t const src = r * W;
                size_t const dst = static_cast<size_t>(y - params->buf_y0) * W;
                if (rgb) ParallelMemcpy(rgb + dst * 3, static_cast<float*>(m_h_rgb) + src * 3, run * W * 12);
                if (hit_id) std::memcpy(hit_id + dst, static_cast<int32_t*>(m_h_id) + src, run * W * 4);
                if (hit_t) std::memcpy(hit_t + dst, static_cast<float*>(m_h_t) + src, run * W * 4);
                r += run;
            }
        }
        if (direct) ML_CUDA_CHECK(cudaStreamSynchronize(sc));
        ML_CUDA_CHECK(cudaStreamSynchronize(s0));
        if (stats) return GetStats(stats);
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_error = e.what();
        return ML_FAIL;
    }
}

} // namespace ML

// ---------------------------------------------------------------------------
// C entry points
// ---------------------------------------------------------------------------
extern "C" {

ML_API_ENTRY srt_scene srtCreateScene(ml_context context, size_t n_triangles,
                                      float const* vertices, float const* albedo, int build_bvh)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr) return nullptr;
    ctx->ClearError();
    try
    {
        return ML::Scene::MakeHandle(
            new ML::Scene(ctx->Device(), n_triangles, vertices, albedo, build_bvh != 0));
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return nullptr;
    }
}

ML_API_ENTRY ml_status srtGetSceneInfo(srt_scene scene, srt_scene_info* info)
{
    if (ML::Scene::FromHandle(scene) == nullptr) return ML_FAIL;
    return ML::Scene::FromHandle(scene)->GetInfo(info);
}

ML_API_ENTRY char* srtGetSceneError(srt_scene scene, char* buffer, size_t buffer_size)
{
    if (ML::Scene::FromHandle(scene) == nullptr)
        return ML::FillBuffer(buffer, buffer_size, "Bad scene handle");
    return ML::Scene::FromHandle(scene)->GetError(buffer, buffer_size);
}

ML_API_ENTRY void srtReleaseScene(srt_scene scene)
{
    delete ML::Scene::FromHandle(scene);
}

ML_API_ENTRY size_t srtGetSceneBvh(srt_scene scene, float* nodes, size_t max_nodes,
                                   int32_t* order, size_t max_order)
{
    if (ML::Scene::FromHandle(scene) == nullptr) return 0;
    return ML::Scene::FromHandle(scene)->GetBvh(nodes, max_nodes, order, max_order);
}

ML_API_ENTRY size_t srtGenCornellBox(float* vertices, float* albedo, srt_camera* camera,
                                     srt_light* light, uint32_t width, uint32_t height)
{
    return ML::GenCornellBox(vertices, albedo, camera, light, width, height);
}

ML_API_ENTRY ml_status srtGenMesh(uint32_t seed, size_t n_triangles, float* vertices,
                                  float* albedo, srt_camera* camera, srt_light* light,
                                  uint32_t width, uint32_t height)
{
    try
    {
        ML::GenMesh(seed, n_triangles, vertices, albedo, camera, light, width, height);
        return ML_OK;
    }
    catch (...)
    {
        return ML_FAIL;
    }
}

ML_API_ENTRY ml_status srtRender(srt_scene scene, srt_camera const* camera,
                                 srt_light const* light, srt_render_params const* params,
                                 float* rgb, int32_t* hit_id, float* hit_t,
                                 srt_render_stats* stats)
{
    if (ML::Scene::FromHandle(scene) == nullptr) return ML_FAIL;
    return ML::Scene::FromHandle(scene)->RenderHost(camera, light, params, rgb, hit_id, hit_t, stats);
}

ML_API_ENTRY ml_status srtRenderDevice(srt_scene scene, srt_camera const* camera,
                                       srt_light const* light, srt_render_params const* params,
                                       void* d_rgb, void* d_hit_id, void* d_hit_t, void* stream)
{
    if (ML::Scene::FromHandle(scene) == nullptr) return ML_FAIL;
    return ML::Scene::FromHandle(scene)->RenderDevice(camera, light, params, d_rgb, d_hit_id, d_hit_t,
                                                      static_cast<cudaStream_t>(stream), true);
}

ML_API_ENTRY ml_status srtRenderDeviceNoAlloc(srt_scene scene, srt_camera const* camera,
                                              srt_light const* light, srt_render_params const* params,
                                              void* d_rgb, void* d_hit_id, void* d_hit_t, void* stream)
{
    if (ML::Scene::FromHandle(scene) == nullptr) return ML_FAIL;
    return ML::Scene::FromHandle(scene)->RenderDeviceNoAlloc(camera, light, params, d_rgb, d_hit_id, d_hit_t,
                                                             static_cast<cudaStream_t>(stream));
}

ML_API_ENTRY ml_status srtGetRenderStats(srt_scene scene, srt_render_stats* stats)
{
    if (ML::Scene::FromHandle(scene) == nullptr) return ML_FAIL;
    return ML::Scene::FromHandle(scene)->GetStats(stats);
}

ML_API_ENTRY ml_status srtRenderToImage(srt_scene scene, srt_camera const* camera,
                                        srt_light const* light, srt_render_params const* params,
                                        ml_image image, srt_render_stats* stats)
{
    ML::Scene* sc = ML::Scene::FromHandle(scene);
    if (sc == nullptr) return ML_FAIL;
    ML::Image* img = ML::Image::FromHandle(image);
    if (img == nullptr) { sc->SetError("Bad image handle"); return ML_FAIL; }
    if (params == nullptr) { sc->SetError("Bad render parameters argument"); return ML_FAIL; }
    ml_image_info const& info = img->Info();
    if (info.dtype != ML_FLOAT32 || info.channels != 3 || info.width != params->width
        || info.height != params->height)
    {
        std::ostringstream os;
        os << "Image " << info.width << " x " << info.height << " x " << info.channels
           << " does not match a FLOAT32 frame " << params->width << " x " << params->height << " x 3";
        sc->SetError(os.str());
        return ML_FAIL;
    }
    if (img->Device()->device != sc->Device()->device)
    {
        sc->SetError("Image and scene live on different devices");
        return ML_FAIL;
    }
    try
    {
        cudaStream_t s = sc->Device()->stream;
        srt_render_params p = *params;
        p.buf_y0 = 0;
        // A partial band must keep the other rows: make the device copy current first.
        bool const full = (p.y0 == 0 && p.y1 == p.height);
        void* d = full ? img->DeviceWrite() : img->DeviceRead(s);
        ml_status st = sc->RenderDevice(camera, light, &p, d, nullptr, nullptr, s, true);
        if (st != ML_OK) return st;
        img->CommitDeviceWrite(s);
        ML_CUDA_CHECK(cudaStreamSynchronize(s));
        if (stats) return sc->GetStats(stats);
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        sc->SetError(e.what());
        return ML_FAIL;
    }
}

ML_API_ENTRY double srtMeasureFp32Peak(ml_context context, int repeats)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr) return 0.0;
    try
    {
        return ML::MeasureFp32Peak(*ctx->Device(), repeats);
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return 0.0;
    }
}

} // extern "C"
