// scene.h -- ML::Scene: triangle set + BVH resident in HBM (BUILD-DEFINED; the
// reference has no scene code).
//
// HBM layout (all float4 = 16 B, 16-byte aligned, SoA so that a tile of
// triangles is three contiguous byte ranges a bulk copy can stage into shared
// memory):
//   tri_v0[i] = (v0.x, v0.y, v0.z, bits(original triangle id))
//   tri_e1[i] = (v1 - v0, 0)        tri_e2[i] = (v2 - v0, 0)
//   tri_n[i]  = (normalize(e1 x e2), 0)             same slot order, filled by a device kernel at upload
//   albedo[id] = (r, g, b, 0)                       indexed by ORIGINAL id
//   nodes[4*k .. 4*k+3]: one 64-byte inner BVH node (two child boxes + refs)
// Triangles are stored in BVH leaf order (identity when no BVH is built) and
// padded with never-hit zero triangles (id -1) to a multiple of kTriTile so the
// brute-force sweep needs no bounds checks.
#pragma once

#include "common.h"
#include "srt_render.h"

#include <cstdint>
#include <vector>

namespace ML {

struct F4 { float x, y, z, w; };

constexpr int kTriTile = 256;      // triangles per shared-memory tile (12 KB)
constexpr int kMaxLeafTris = 8;    // largest BVH leaf the encoding is used with
constexpr int kDefaultLeafTris = 2; // BVH leaf size
constexpr int kLeafCountBits = 4;  // child ref: ~ref = (first << 4) | (count-1)
constexpr int kBvhStackDepth = 64;

// 64-byte inner node as stored in HBM:
//   n0 = (c0.min.x, c1.min.x, c0.max.x, c1.max.x)
//   n1 = (c0.min.y, c1.min.y, c0.max.y, c1.max.y)
//   n2 = (c0.min.z, c1.min.z, c0.max.z, c1.max.z)
//   n3 = (bits(ref0), bits(ref1), 0, 0)
struct BvhNode { F4 n0, n1, n2, n3; };

struct BvhBuildResult
{
    std::vector<BvhNode> nodes;     // node 0 is the root (always an inner node)
    std::vector<int32_t> order;     // leaf order -> original triangle id
    int max_depth = 0;
};

// Deterministic binned-SAH top-down build over triangle bounds (host).
BvhBuildResult BuildBvh(size_t n, float const* vertices /*9 per tri*/);

struct DeviceSceneView
{
    F4 const* tri_v0;
    F4 const* tri_e1;
    F4 const* tri_e2;
    F4 const* tri_n;        // per slot: geometric unit normal normalize(e1 x e2), computed ONCE on the device at
                            // upload with the exact operation sequence of docs/RENDER_SPEC.md section 4
    F4 const* albedo;
    F4 const* nodes;
    uint32_t n_tris;        // real triangles
    uint32_t n_tris_padded; // multiple of kTriTile
};

// Screen-space candidate grid for primary rays (BUILD-DEFINED, docs/RENDER_SPEC.md section 8): per
// 8x8-pixel tile the triangles whose conservatively dilated screen bounding box touches the tile.
// All primary rays leave one point, so a ray through pixel (x, y) can only hit triangles whose
// projection covers that pixel: the closest hit over the tile's (pixel-filtered) list equals the
// closest hit over the whole scene, and the lexicographic (t, id) rule makes the test order
// irrelevant.  Rebuilt on the device for every render call (three small kernels); a frame the grid
// cannot serve (flags.ok == 0) and tiles with long lists fall back to the BVH.
constexpr uint32_t kGridTile = 8;        // pixels; equals the render kernels' 8-row blocks
constexpr uint32_t kGridScanCells = 1024; // cells per block of the grid build's prefix sum (sizes GridBuild::block_sums)
constexpr uint32_t kSppMaxCand = 256;    // sample-parallel kernel: candidates (tile list + global list) it can stage; a tile
                                         // with more walks the BVH instead (one byte indexes a candidate)
constexpr uint32_t kGridMaxGlobal = 32;  // triangles that cannot be projected safely (near or crossing the
                                         // camera plane, or huge on screen): tested by every ray
// ticket: blocks of the build's block-sum launch that have finished (the last one scans the sums); zeroed with the rest
struct GridFlags { uint32_t ok; uint32_t n_global; uint32_t total; uint32_t ticket; };
struct GridView
{
    uint32_t const* tile_off;   // [tiles + 1] exclusive prefix of the list lengths
    uint32_t const* items;      // triangle slots, tile after tile
    F4 const* bbox;             // per slot: dilated screen box (x0, y0, x1, y1) in pixels; x0 > x1: not binned
    F4 const* edges;            // per slot 3 x (nx, ny, c, *): the projected triangle's edge functions with the
                                // conservative margin folded into c -- a query at q can only hit the triangle if
                                // nx*q.x + ny*q.y + c >= 0 for all three; edges[3*slot].w = nearest depth (light grid)
    uint32_t const* global;     // [kGridMaxGlobal]
    GridFlags const* flags;
    uint32_t tiles_w, tiles_h;
    uint32_t enabled;           // host decision for this launch
    // Heaviest-first work list (screen grid): the owned tiles sorted into kGridBuckets classes of estimated cost,
    // order[b * order_stride + i] = i-th item of class b (bits 0..23 tile; bits 28..31 = columns, 0 = the whole
    // tile, else that many of its eight pixel columns starting at bits 24..27), bucket_cnt[b] = items in class b.
    // A launch that covers exactly the band the grid was built for runs persistent blocks over this list, so the
    // long items start first and the launch drains on short ones.
    uint32_t const* order;
    uint32_t const* bucket_cnt;
    uint32_t order_stride;
    uint32_t* tile_time;        // [tiles] ns per tile of this frame, for the next frame's list (NULL: not recorded)
};
constexpr uint32_t kGridBuckets = 8;

// The same structure seen from the point light, for shadow rays: every shadow segment ends at the
// light, so the segment from a shading point S projects to ONE point of a perspective grid centred
// on the light, and only triangles whose (dilated) projection covers that cell can occlude it.
// Cells are single grid units (tile = 1); a query is  c = minv * (S - light),  cell = (cx + kx*c.x/c.z,
// cy - ky*c.y/c.z).  S behind the light's projection plane (c.z <= 0) walks the BVH instead.
struct LightGridView
{
    GridView grid;
    float minv[9];
    float kx, ky, cx, cy;
    uint32_t max_walk;      // list entries a shadow ray tests before it hands the query to the BVH (SRT_LIGHT_WALK)
};

class Scene
{
public:
    static srt_scene MakeHandle(Scene* s) { return reinterpret_cast<srt_scene>(s); }
    static Scene* FromHandle(srt_scene s) { return reinterpret_cast<Scene*>(s); }

    Scene(DeviceStatePtr device, size_t n_triangles, float const* vertices,
          float const* albedo, bool build_bvh, BvhBuildResult const* prebuilt = nullptr);
    ~Scene();
    Scene(Scene const&) = delete;
    Scene& operator=(Scene const&) = delete;

    ml_status GetInfo(srt_scene_info* info) const;
    size_t GetBvh(float* nodes, size_t max_nodes, int32_t* order, size_t max_order) const;

    // Enqueues the render on `stream`; with `timed` the kernel is bracketed by
    // the context's events (read back by GetStats).
    // `signal`: device-side completion flags for multi-GPU frames.  The band is rendered as n_chunks launches
    // (ChunkLayout); chunk k's last block stores `value` to d_flags[k * chunk_stride] (RenderLaunch::d_signal).
    struct SignalSpec { uint32_t* d_flags; uint32_t chunk_stride; uint32_t n_chunks; uint32_t value; };
    ml_status RenderDevice(srt_camera const* camera, srt_light const* light,
                           srt_render_params const* params, void* d_rgb, void* d_hit_id,
                           void* d_hit_t, cudaStream_t stream, bool timed, SignalSpec const* signal = nullptr);
    // srtRenderDeviceNoAlloc: RenderDevice without any allocation (static scratch, no events, no light grid).
    ml_status RenderDeviceNoAlloc(srt_camera const* camera, srt_light const* light, srt_render_params const* params,
                                  void* d_rgb, void* d_hit_id, void* d_hit_t, cudaStream_t stream);
    ml_status RenderHost(srt_camera const* camera, srt_light const* light,
                         srt_render_params const* params, float* rgb, int32_t* hit_id,
                         float* hit_t, srt_render_stats* stats);
    ml_status GetStats(srt_render_stats* stats);

    DeviceStatePtr const& Device() const { return m_device; }
    char* GetError(char* buffer, size_t n) const { return FillBuffer(buffer, n, m_error); }
    void SetError(std::string const& m) { m_error = m; }
    bool HasBvh() const { return !m_nodes_host.empty(); }

private:
    bool Validate(srt_camera const* camera, srt_light const* light,
                  srt_render_params const* params);
    void EnqueueBand(srt_camera const* camera, srt_light const* light, srt_render_params const* params,
                     void* d_rgb, void* d_hit_id, void* d_hit_t, cudaStream_t stream,
                     uint32_t* d_signal, uint32_t signal_value, uint32_t ticket, bool whole_band);
    void EnsureStaging(size_t pixels);
    void FreeStaging();
    // Builds the primary-ray grid for this camera / frame on `stream` (if the render will use it).
    void PrepareGrid(srt_camera const* camera, srt_render_params const* params, cudaStream_t stream);
    void PrepareLightGrid(srt_light const* light, srt_render_params const* params, cudaStream_t stream);
    void FreeGrid();

    DeviceStatePtr m_device;
    srt_scene_info m_info{};
    DeviceSceneView m_view{};
    void* m_d_tris = nullptr;     // one allocation: v0 | e1 | e2
    void* m_d_albedo = nullptr;
    void* m_d_nodes = nullptr;
    unsigned long long* m_d_counters = nullptr; // per-frame scratch: primary, shadow, box, tri | tickets | grid flags
    std::vector<BvhNode> m_nodes_host;
    std::vector<int32_t> m_order_host;
    // staging for RenderHost (device band buffers + pinned host twins)
    void* m_d_rgb = nullptr; void* m_d_id = nullptr; void* m_d_t = nullptr;
    void* m_h_rgb = nullptr; void* m_h_id = nullptr; void* m_h_t = nullptr;
    size_t m_stage_pixels = 0;
    // primary-ray grid (device)
    GridView m_grid{};
    void* m_d_grid_tiles = nullptr;   // tile_cnt | tile_off | cursor
    void* m_d_grid_items = nullptr;
    void* m_d_grid_bbox = nullptr;    // bbox | edges
    void* m_d_grid_misc = nullptr;    // GridFlags | global list
    void* m_d_grid_order = nullptr;   // m_grid_tiles tile times | kGridBuckets x 8 x m_grid_tiles work items
    uint32_t m_grid_tiles = 0;
    uint32_t m_grid_cap = 0;
    // shadow-ray grid (device)
    LightGridView m_lgrid{};
    void* m_d_lgrid_tiles = nullptr;
    void* m_d_lgrid_items = nullptr;
    void* m_d_lgrid_bbox = nullptr;
    void* m_d_lgrid_misc = nullptr;
    uint32_t m_lgrid_tiles = 0;
    uint32_t m_lgrid_cap = 0;
    bool m_lgrid_valid = false;       // the light grid on the device was built for m_lgrid_light
    float m_lgrid_light[3] = {0.f, 0.f, 0.f};
    float m_pad = 0.f;                // world-space slack of the culling structures (as the BVH boxes)
    cudaEvent_t m_grid_free = nullptr; // the last render that read the grid has finished
    cudaEvent_t m_lgrid_built = nullptr; // the cached light grid is complete
    cudaEvent_t m_lgrid_ev0 = nullptr;   // start of the light grid build (timing)
    cudaEvent_t m_ev_grid_done = nullptr; // grids built, render kernel next (srt_render_stats::grid_build_ms)
    cudaEvent_t m_ev_fork = nullptr, m_ev_join = nullptr; // chunked signalled render: helper stream fork / join
    bool m_timed_grid = false;
    std::vector<cudaEvent_t> m_chunk_events; // RenderHost pipeline
    cudaStream_t m_timed_stream = nullptr;
    bool m_timed_pending = false;
    std::string m_error;
};

// Scene generators (host, scenes.cpp).
size_t GenCornellBox(float* vertices, float* albedo, srt_camera* camera, srt_light* light,
                     uint32_t width, uint32_t height);
void GenMesh(uint32_t seed, size_t n_triangles, float* vertices, float* albedo,
             srt_camera* camera, srt_light* light, uint32_t width, uint32_t height);

// Kernel launchers (render_kernels.cu).
struct RenderLaunch
{
    DeviceSceneView scene;
    srt_camera camera;
    srt_light light;
    srt_render_params params;
    float inv2w, inv2h, inv_spp;
    float* d_rgb;
    int32_t* d_hit_id;
    float* d_hit_t;
    unsigned long long* d_counters;
    // Rows this launch owns: 8-row blocks k = 0..n_row_blocks-1 start at frame
    // row  row_origin + k * row_block_step  (step 8: the whole band; step 8*N:
    // every N-th block, interleaved multi-GPU sharding).  Rows >= y1 are clipped.
    uint32_t row_origin;
    uint32_t row_block_step;
    uint32_t n_row_blocks;
    uint32_t ordered; // 1 (sample-parallel kernel, one pixel per warp pass): persistent blocks pull work items from
                      // grid.order (heaviest class first, heavy tiles split in four) through the cursor *d_work
    uint32_t* d_work; // zeroed by the per-frame scratch memset
    uint32_t hints;  // 1: rays first test the triangle the lane's previous ray ended on (SRT_BVH_HINTS=0 disables)
    // Completion signal fused into the render kernel (multi-GPU frames): when d_signal != NULL the LAST block
    // of the launch to finish (atomic ticket in d_ticket, reset for the next launch) stores signal_value to
    // *d_signal with system-scope release ordering, after every block's pixels (possibly peer stores over
    // NVLink into another GPU's frame) have been fenced.  d_signal may itself be peer memory.
    // SRT_BLOCK_TRACE=1 (diagnostic): 5 x u64 {min block start, max block start, max block end, sum of block
    // durations, max block duration} in ns of %globaltimer; NULL otherwise
    unsigned long long* d_trace;
    uint32_t* d_signal;
    uint32_t* d_ticket;
    uint32_t signal_value;
    GridView grid;
    LightGridView lgrid;
};
void LaunchRender(RenderLaunch const& launch, int sm_count, cudaStream_t stream);
// Which candidate grids the kernel LaunchRender picks for these parameters reads: bit 0 screen grid, bit 1 light grid.
int RenderUsesGrid(srt_render_params const& params);
void ChunkLayout(srt_render_params const& params, uint32_t requested, uint32_t& chunk_rows, uint32_t& n_chunks);
struct GridBuild
{
    DeviceSceneView scene;
    float origin[3];
    float minv[9];          // rows of the inverse of [right up forward]: c = minv * (P - origin)
    float half_w, half_h;
    float pad;              // world-space slack
    uint32_t width, height, tiles_w, tiles_h, cap;
    uint32_t tile;          // grid units per tile: 8 (pixels, camera grid) or 1 (light grid)
    uint32_t max_tiles_per_tri; // a triangle whose box spans more tiles goes to the global list
    // tile rows this call renders: [ty_begin, ty_end), and with ty_stride > 1 only rows ty % ty_stride == ty_phase
    // (interleaved multi-GPU sharding); other rows get no entries.  The light grid covers every row.
    uint32_t ty_begin, ty_end, ty_stride, ty_phase;
    float lo_bias;          // 1: queries are closed unit squares [x, x+1] (pixels); 0: points
    uint32_t* tile_cnt; uint32_t* tile_off; uint32_t* cursor; uint32_t* items;
    uint32_t* block_sums;   // [tiles / kGridScanCells + 1] scratch of the prefix sum
    uint32_t* tile_time;    // [tiles] or NULL: ns each tile took in the previous frame (render kernel adds, build clears)
    uint32_t* order;        // [kGridBuckets * order_stride] or NULL (light grid, no-alloc path): see GridView::order
    uint32_t* bucket_cnt;   // [kGridBuckets], zeroed by the caller's per-frame scratch memset
    uint32_t order_stride;
    F4* bbox; F4* edges; uint32_t* global; GridFlags* flags;
    float half_extent;      // half size of a query in grid units: 0.5 (pixel squares) or 0 (points)
    uint32_t flags_prezeroed; // 1: the caller's per-frame scratch memset already cleared *flags
};
void BuildScreenGrid(GridBuild const& build, cudaStream_t stream);
// Consumer / producer halves of the device-side completion flags (render_kernels.cu).
void LaunchWaitFlags(uint32_t const* d_flags, uint32_t count, uint32_t stride, uint32_t value,
                     uint32_t timeout_ms, uint32_t* timeouts_mapped, cudaStream_t stream);
void LaunchWriteFlag(uint32_t* d_flag, uint32_t value, cudaStream_t stream);
// tri_n[i] = normalize(tri_e1[i] x tri_e2[i]) for all padded slots (scene upload).
void LaunchTriNormals(F4 const* e1, F4 const* e2, F4* n, uint32_t count, cudaStream_t stream);
// Sorts every cell's list of a (light) grid by the triangles' nearest depth (edges[3*slot].w), ascending.
void LaunchSortCellsByDepth(GridBuild const& build, cudaStream_t stream);
// srtRenderDeviceNoAlloc (render_kernels.cu): grid build + render with every scratch pointer in static device memory.
void RenderNoAllocScratch(int device, RenderLaunch launch, GridBuild build, bool use_grid, int sm_count,
                          cudaStream_t stream);
double MeasureFp32Peak(DeviceState const& device, int repeats);

} // namespace ML
