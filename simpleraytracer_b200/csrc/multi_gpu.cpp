// multi_gpu.cpp -- one frame on N GPUs (BUILD-DEFINED, the reference is single-device).
//
// Exchange = peer stores + device-side flags.  Every GPU renders its interleaved 8-row blocks straight into
// the frame buffer of GPU 0 (NVLink / NVSwitch peer mapping); the render kernel's last block publishes a
// completion flag next to it (render_kernels.cu flush_counters) and GPU 0 waits for the flags with a spin
// kernel on its own stream (wait_flags_kernel).  No gather pass, no host round trip, no collective.
//
// Two drivers share this machinery:
//  * one process per GPU (bench.py under torchrun): srtRenderDeviceSignal on every rank, frame and flags
//    IPC-mapped from rank 0 (mlxIpcExportImage / mlxIpcExportPointer / mlxIpcOpen), srtStreamWaitFlags or
//    srtGatherFrameHost on rank 0;
//  * one process, N GPUs (SceneGroup below, srtRenderGroup): cudaDeviceEnablePeerAccess, one persistent host
//    thread per device, no torch, no NCCL.
#include "scene.h"

#include "context.h"
#include "model_runner_ext.h"

#include <algorithm>
#include <condition_variable>
#include <mutex>
#include <thread>

namespace ML {
namespace {

bool PageLocked(void const* ptr)
{
    cudaPointerAttributes at{};
    if (cudaPointerGetAttributes(&at, ptr) != cudaSuccess) { cudaGetLastError(); return false; }
    return at.type == cudaMemoryTypeHost;
}

// GPU 0's half of a host-buffer frame (srtGatherFrameHost).  Runs on the device's copy_out stream.
void GatherFrameHost(DeviceState& dev, srt_render_params const& p, void const* d_rgb, uint32_t const* d_flags,
                     uint32_t n_ranks, uint32_t requested_chunks, uint32_t value, float* rgb_host,
                     uint32_t timeout_ms)
{
    dev.MakeCurrent();
    cudaStream_t const sc = dev.copy_out;
    uint32_t chunk_rows = 0, n_chunks = 0;
    ChunkLayout(p, requested_chunks, chunk_rows, n_chunks);
    if (n_chunks > 16) throw std::runtime_error("srtGatherFrameHost: more than 16 chunks");
    size_t const row_bytes = static_cast<size_t>(p.width) * 12u;
    bool const direct = PageLocked(rgb_host);
    size_t const band_bytes = static_cast<size_t>(p.y1 - p.y0) * row_bytes;
    if (!direct && dev.gather_stage_bytes < band_bytes)
    {
        if (dev.gather_stage) ML_CUDA_CHECK(cudaFreeHost(dev.gather_stage));
        dev.gather_stage = nullptr;
        dev.gather_stage_bytes = 0;
        ML_CUDA_CHECK(cudaMallocHost(&dev.gather_stage, band_bytes));
        dev.gather_stage_bytes = band_bytes;
    }
    uint32_t const before = *dev.wait_timeouts;
    char const* const src = static_cast<char const*>(d_rgb) + static_cast<size_t>(p.y0 - p.buf_y0) * row_bytes;
    char* const dst = reinterpret_cast<char*>(rgb_host) + static_cast<size_t>(p.y0 - p.buf_y0) * row_bytes;
    // Every flag chunk is copied as up to `pieces` transfers, so that with a pageable destination the host copy of
    // piece k runs while piece k + 1 crosses PCIe -- also when the frame is ONE flag chunk (a work-list render
    // finishes rows in no particular order and signals once).  At most 8 pieces for a multi-chunk frame: a piece must
    // stay large enough for the copy pool (16 pieces of a 1080p frame are 1.5 MB each, which one thread copied alone
    // at ~11 GB/s: 2.2 ms of host copy behind a 0.6 ms render at 8 GPUs).
    uint32_t const pieces = direct ? 1u : std::max(1u, std::min(4u, 8u / n_chunks));
    struct Piece { size_t r0, r1; };
    Piece piece[16];
    uint32_t n_pieces = 0;
    for (uint32_t k = 0; k < n_chunks; k++)
    {
        size_t const c0 = static_cast<size_t>(k) * chunk_rows, c1 = std::min<size_t>(p.y1 - p.y0, c0 + chunk_rows);
        LaunchWaitFlags(d_flags + static_cast<size_t>(k) * n_ranks, n_ranks, 1u, value,
                        timeout_ms ? timeout_ms : 2000u, dev.wait_timeouts, sc);
        size_t const step = (c1 - c0 + pieces - 1) / pieces;
        for (size_t r0 = c0; r0 < c1; r0 += step)
        {
            size_t const r1 = std::min(c1, r0 + step);
            char* const h = direct ? dst : static_cast<char*>(dev.gather_stage);
            ML_CUDA_CHECK(cudaMemcpyAsync(h + r0 * row_bytes, src + r0 * row_bytes, (r1 - r0) * row_bytes,
                                          cudaMemcpyDeviceToHost, sc));
            if (!direct)
            {
                if (dev.gather_events[n_pieces] == nullptr)
                    ML_CUDA_CHECK(cudaEventCreateWithFlags(&dev.gather_events[n_pieces], cudaEventDisableTiming));
                ML_CUDA_CHECK(cudaEventRecord(dev.gather_events[n_pieces], sc));
            }
            piece[n_pieces++] = Piece{r0, r1};
        }
    }
    // pageable destination: move each piece out of the staging buffer as it lands, while later ones are in flight
    for (uint32_t k = 0; k < n_pieces && !direct; k++)
    {
        ML_CUDA_CHECK(cudaEventSynchronize(dev.gather_events[k]));
        ParallelMemcpy(dst + piece[k].r0 * row_bytes, static_cast<char*>(dev.gather_stage) + piece[k].r0 * row_bytes,
                       (piece[k].r1 - piece[k].r0) * row_bytes);
    }
    ML_CUDA_CHECK(cudaStreamSynchronize(sc));
    if (*dev.wait_timeouts != before)
        throw std::runtime_error("multi-GPU frame: a completion flag did not arrive within the timeout "
                                 "(a rank did not render, or peer stores are not reaching GPU 0)");
}

} // namespace
} // namespace ML

extern "C" {

ML_API_ENTRY ml_status srtChunkLayout(srt_render_params const* params, uint32_t requested_chunks,
                                      uint32_t* chunk_rows, uint32_t* n_chunks)
{
    if (params == nullptr || params->y0 > params->y1) return ML_FAIL;
    uint32_t rows = 0, n = 0;
    ML::ChunkLayout(*params, requested_chunks, rows, n);
    if (chunk_rows) *chunk_rows = rows;
    if (n_chunks) *n_chunks = n;
    return ML_OK;
}

ML_API_ENTRY ml_status srtRenderDeviceSignal(srt_scene scene, srt_camera const* camera,
                                             srt_light const* light, srt_render_params const* params,
                                             void* d_rgb, void* d_hit_id, void* d_hit_t, void* stream,
                                             void* d_flags, uint32_t chunk_stride, uint32_t n_chunks,
                                             uint32_t value)
{
    ML::Scene* sc = ML::Scene::FromHandle(scene);
    if (sc == nullptr) return ML_FAIL;
    if (d_flags == nullptr) { sc->SetError("Bad flags argument"); return ML_FAIL; }
    ML::Scene::SignalSpec sig{static_cast<uint32_t*>(d_flags), chunk_stride, n_chunks ? n_chunks : 1u, value};
    return sc->RenderDevice(camera, light, params, d_rgb, d_hit_id, d_hit_t, static_cast<cudaStream_t>(stream),
                            true, &sig);
}

ML_API_ENTRY ml_status srtStreamWaitFlags(ml_context context, void* stream, void const* d_flags,
                                          uint32_t count, uint32_t stride, uint32_t value, uint32_t timeout_ms)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr) return ML_FAIL;
    ctx->ClearError();
    try
    {
        if (d_flags == nullptr) throw std::runtime_error("Bad flags argument");
        ctx->Device()->MakeCurrent();
        ML::LaunchWaitFlags(static_cast<uint32_t const*>(d_flags), count, stride ? stride : 1u, value,
                            timeout_ms ? timeout_ms : 2000u, ctx->Device()->wait_timeouts,
                            stream ? static_cast<cudaStream_t>(stream) : ctx->Device()->stream);
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return ML_FAIL;
    }
}

ML_API_ENTRY ml_status srtStreamWriteFlag(ml_context context, void* stream, void* d_flag, uint32_t value)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr) return ML_FAIL;
    ctx->ClearError();
    try
    {
        if (d_flag == nullptr) throw std::runtime_error("Bad flag argument");
        ctx->Device()->MakeCurrent();
        ML::LaunchWriteFlag(static_cast<uint32_t*>(d_flag), value,
                            stream ? static_cast<cudaStream_t>(stream) : ctx->Device()->stream);
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return ML_FAIL;
    }
}

ML_API_ENTRY uint32_t srtFlagWaitTimeouts(ml_context context)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    return ctx ? *ctx->Device()->wait_timeouts : 0u;
}

ML_API_ENTRY ml_status srtGatherFrameHost(ml_context context, srt_render_params const* params,
                                          void const* d_rgb, void const* d_flags, uint32_t n_ranks,
                                          uint32_t n_chunks, uint32_t value, float* rgb_host,
                                          uint32_t timeout_ms)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr) return ML_FAIL;
    ctx->ClearError();
    try
    {
        if (params == nullptr || d_rgb == nullptr || d_flags == nullptr || rgb_host == nullptr || n_ranks == 0
            || params->y0 > params->y1 || params->buf_y0 > params->y0)
            throw std::runtime_error("Bad gather argument");
        ML::GatherFrameHost(*ctx->Device(), *params, d_rgb, static_cast<uint32_t const*>(d_flags), n_ranks,
                            n_chunks, value, rgb_host, timeout_ms);
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return ML_FAIL;
    }
}

ML_API_ENTRY void* mlxDeviceAlloc(ml_context context, size_t bytes)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr) return nullptr;
    ctx->ClearError();
    try
    {
        ctx->Device()->MakeCurrent();
        // whole 2 MiB granules: an exported (IPC) buffer then shares its physical block with nothing else
        size_t const granule = size_t(2) << 20;
        size_t const n = (std::max<size_t>(bytes, 1) + granule - 1) / granule * granule;
        void* p = nullptr;
        ML_CUDA_CHECK(cudaMalloc(&p, n));
        ML_CUDA_CHECK(cudaMemset(p, 0, n));
        return p;
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return nullptr;
    }
}

ML_API_ENTRY void mlxDeviceFree(ml_context context, void* device_pointer)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr || device_pointer == nullptr) return;
    cudaSetDevice(ctx->Device()->device);
    cudaFree(device_pointer);
}

ML_API_ENTRY ml_status mlxIpcExportPointer(ml_context context, void* device_pointer, void* handle_out)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr) return ML_FAIL;
    ctx->ClearError();
    try
    {
        if (device_pointer == nullptr || handle_out == nullptr) throw std::runtime_error("Bad pointer argument");
        ctx->Device()->MakeCurrent();
        cudaIpcMemHandle_t h;
        ML_CUDA_CHECK(cudaIpcGetMemHandle(&h, device_pointer));
        std::memcpy(handle_out, &h, sizeof(h));
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return ML_FAIL;
    }
}

} // extern "C"

// ---------------------------------------------------------------------------
// SceneGroup: one process, N GPUs
// ---------------------------------------------------------------------------
namespace ML {

class SceneGroup
{
public:
    static srt_scene_group MakeHandle(SceneGroup* g) { return reinterpret_cast<srt_scene_group>(g); }
    static SceneGroup* FromHandle(srt_scene_group g) { return reinterpret_cast<SceneGroup*>(g); }

    SceneGroup(int const* devices, uint32_t n, size_t n_triangles, float const* vertices, float const* albedo,
               bool build_bvh);
    ~SceneGroup();
    SceneGroup(SceneGroup const&) = delete;
    SceneGroup& operator=(SceneGroup const&) = delete;

    ml_status Render(srt_camera const* camera, srt_light const* light, srt_render_params const* params,
                     float* rgb_host, srt_render_stats* stats);
    void* Frame() const { return m_d_frame; }
    uint32_t Size() const { return static_cast<uint32_t>(m_members.size()); }
    char* GetError(char* buffer, size_t n) const { return FillBuffer(buffer, n, m_error); }

private:
    struct Member
    {
        DeviceStatePtr device;
        std::unique_ptr<Scene> scene;
        std::thread thread;
        ml_status status = ML_OK;
        cudaEvent_t done = nullptr;   // same-GPU replicas only: their launches have completed
    };
    struct Job
    {
        srt_camera camera;
        srt_light light;
        srt_render_params params;
        uint32_t n_chunks;
        uint32_t value;
    };

    void Worker(uint32_t rank);
    void EnsureFrame(size_t bytes);

    std::vector<Member> m_members;
    void* m_d_frame = nullptr;       // on m_members[0]'s GPU
    size_t m_frame_bytes = 0;
    uint32_t* m_d_flags = nullptr;   // [16 chunks][n ranks], on m_members[0]'s GPU
    uint32_t m_epoch = 0;
    cudaEvent_t m_ev_end = nullptr;  // GPU 0: every flag seen (and, with a host frame, the last chunk copied)

    std::mutex m_mutex;
    std::condition_variable m_cv_job, m_cv_done;
    uint64_t m_job_id = 0;
    uint32_t m_pending = 0;
    bool m_quit = false;
    Job m_job{};
    std::string m_error;
};

SceneGroup::SceneGroup(int const* devices, uint32_t n, size_t n_triangles, float const* vertices,
                       float const* albedo, bool build_bvh)
{
    if (devices == nullptr || n == 0) throw std::runtime_error("Bad device list");
    if (n > 64) throw std::runtime_error("Too many devices");
    if (n_triangles > 0 && vertices == nullptr) throw std::runtime_error("Bad vertices argument");
    BvhBuildResult bvh;
    if (build_bvh) bvh = BuildBvh(n_triangles, vertices);     // once; every replica uploads the same tree
    m_members.resize(n);
    try
    {
        for (uint32_t r = 0; r < n; r++)
        {
            if (devices[r] < 0) throw std::runtime_error("negative device ordinal");
            m_members[r].device = std::make_shared<DeviceState>(devices[r]);
            m_members[r].scene = std::make_unique<Scene>(m_members[r].device, n_triangles, vertices, albedo,
                                                         build_bvh, build_bvh ? &bvh : nullptr);
        }
        int const owner = m_members[0].device->device;
        for (uint32_t r = 1; r < n; r++)
        {
            int const d = m_members[r].device->device;
            if (d == owner) continue;      // two replicas on one GPU (tests): same address space already
            int can = 0;
            ML_CUDA_CHECK(cudaDeviceCanAccessPeer(&can, d, owner));
            if (!can)
                throw std::runtime_error("GPU " + std::to_string(d) + " has no peer access to GPU "
                                         + std::to_string(owner) + " (the frame owner)");
            ML_CUDA_CHECK(cudaSetDevice(d));
            cudaError_t const e = cudaDeviceEnablePeerAccess(owner, 0);
            if (e == cudaErrorPeerAccessAlreadyEnabled) cudaGetLastError();
            else ML_CUDA_CHECK(e);
        }
        m_members[0].device->MakeCurrent();
        ML_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&m_d_flags), size_t(2) << 20));
        ML_CUDA_CHECK(cudaMemset(m_d_flags, 0, size_t(2) << 20));
        ML_CUDA_CHECK(cudaEventCreate(&m_ev_end));
    }
    catch (...)
    {
        if (m_d_flags) { cudaSetDevice(m_members[0].device->device); cudaFree(m_d_flags); }
        throw;
    }
    for (uint32_t r = 0; r < n; r++) m_members[r].thread = std::thread(&SceneGroup::Worker, this, r);
}

SceneGroup::~SceneGroup()
{
    {
        std::lock_guard<std::mutex> lock(m_mutex);
        m_quit = true;
    }
    m_cv_job.notify_all();
    for (Member& m : m_members)
        if (m.thread.joinable()) m.thread.join();
    for (Member& m : m_members)
    {
        cudaSetDevice(m.device->device);
        cudaStreamSynchronize(m.device->stream);
        if (m.done) cudaEventDestroy(m.done);
        m.scene.reset();
    }
    cudaSetDevice(m_members[0].device->device);
    cudaStreamSynchronize(m_members[0].device->copy_out);
    if (m_ev_end) cudaEventDestroy(m_ev_end);
    cudaFree(m_d_frame);
    cudaFree(m_d_flags);
}

// Device r's host thread: enqueues that device's share of every frame (grid build + interleaved row blocks,
// peer stores into GPU 0's frame, completion flags).  Launch overhead of the N devices runs in parallel.
void SceneGroup::Worker(uint32_t rank)
{
    uint64_t seen = 0;
    while (true)
    {
        Job job;
        {
            std::unique_lock<std::mutex> lock(m_mutex);
            m_cv_job.wait(lock, [&] { return m_quit || m_job_id != seen; });
            if (m_quit) return;
            seen = m_job_id;
            job = m_job;
        }
        Member& me = m_members[rank];
        uint32_t const n = static_cast<uint32_t>(m_members.size());
        srt_render_params p = job.params;
        p.y0 = 0; p.y1 = p.height; p.buf_y0 = 0;
        p.row_block_stride = n > 1 ? n : 0;
        p.row_block_phase = n > 1 ? rank : 0;
        Scene::SignalSpec sig{m_d_flags + rank, n, job.n_chunks, job.value};
        me.status = me.scene->RenderDevice(&job.camera, &job.light, &p, m_d_frame, nullptr, nullptr,
                                           me.device->stream, true, &sig);
        {
            std::lock_guard<std::mutex> lock(m_mutex);
            m_pending--;
        }
        m_cv_done.notify_all();
    }
}

void SceneGroup::EnsureFrame(size_t bytes)
{
    if (bytes <= m_frame_bytes) return;
    for (Member& m : m_members)
    {
        m.device->MakeCurrent();
        ML_CUDA_CHECK(cudaStreamSynchronize(m.device->stream));
    }
    m_members[0].device->MakeCurrent();
    ML_CUDA_CHECK(cudaFree(m_d_frame));
    m_d_frame = nullptr;
    m_frame_bytes = 0;
    ML_CUDA_CHECK(cudaMalloc(&m_d_frame, bytes));
    m_frame_bytes = bytes;
}

ml_status SceneGroup::Render(srt_camera const* camera, srt_light const* light, srt_render_params const* params,
                             float* rgb_host, srt_render_stats* stats)
{
    m_error.clear();
    try
    {
        if (camera == nullptr || light == nullptr || params == nullptr) throw std::runtime_error("Bad render argument");
        if (params->width == 0 || params->height == 0) throw std::runtime_error("Unspecified frame dimension");
        DeviceState& dev0 = *m_members[0].device;
        uint32_t const n = Size();
        EnsureFrame(static_cast<size_t>(params->width) * params->height * 12u);
        srt_render_params whole = *params;
        whole.y0 = 0; whole.y1 = whole.height; whole.buf_y0 = 0;
        whole.row_block_stride = n > 1 ? n : 0;
        whole.row_block_phase = 0;
        int const env_chunks = std::getenv("SRT_GROUP_CHUNKS") ? std::atoi(std::getenv("SRT_GROUP_CHUNKS")) : 0;   // per call: tests flip it
        // one flag chunk: the work-list render is the fastest and the gather pipelines the copy in pieces behind it
        uint32_t want = 1u;
        if (env_chunks > 0) want = static_cast<uint32_t>(std::min(env_chunks, 16));
        uint32_t chunk_rows = 0, n_chunks = 0;
        ChunkLayout(whole, want, chunk_rows, n_chunks);
        uint32_t const value = ++m_epoch;
        {
            std::lock_guard<std::mutex> lock(m_mutex);
            m_job = Job{*camera, *light, *params, n_chunks, value};
            m_pending = n;
            m_job_id++;
        }
        m_cv_job.notify_all();
        // GPU 0's consumer side is enqueued AFTER every worker has issued its launches: a spinning wait kernel
        // must never sit in a hardware work queue ahead of a kernel it is waiting for (streams can share a
        // queue -- CUDA_DEVICE_MAX_CONNECTIONS -- and replicas may share one GPU).
        {
            std::unique_lock<std::mutex> lock(m_mutex);
            m_cv_done.wait(lock, [&] { return m_pending == 0; });
        }
        // Replicas that share the owner's GPU (tests on a one-GPU box) are ordered by a stream dependency instead:
        // kernels that spin on one another must never be separate launches on ONE GPU (nothing guarantees that
        // they run at the same time), so the wait kernel there only ever finds flags that are already set.
        bool shared_gpu = false;
        for (uint32_t r = 1; r < n; r++) shared_gpu = shared_gpu || m_members[r].device->device == dev0.device;
        for (uint32_t r = 0; r < n && shared_gpu; r++)
        {
            Member& m = m_members[r];
            if (m.device->device != dev0.device || m.status != ML_OK) continue;
            m.device->MakeCurrent();
            if (m.done == nullptr) ML_CUDA_CHECK(cudaEventCreateWithFlags(&m.done, cudaEventDisableTiming));
            ML_CUDA_CHECK(cudaEventRecord(m.done, m.device->stream));
            ML_CUDA_CHECK(cudaStreamWaitEvent(dev0.copy_out, m.done, 0));
        }
        dev0.MakeCurrent();
        uint32_t const before = *dev0.wait_timeouts;
        std::string gather_error;
        try
        {
            if (rgb_host)
                GatherFrameHost(dev0, whole, m_d_frame, m_d_flags, n, n_chunks, value, rgb_host, 0);
            else
                LaunchWaitFlags(m_d_flags, n * n_chunks, 1u, value, 2000u, dev0.wait_timeouts, dev0.copy_out);
            ML_CUDA_CHECK(cudaEventRecord(m_ev_end, dev0.copy_out));
            ML_CUDA_CHECK(cudaStreamSynchronize(dev0.copy_out));
        }
        catch (std::exception const& e)
        {
            gather_error = e.what();     // a worker's own error (below) explains a missing flag better
        }
        for (uint32_t r = 0; r < n; r++)
            if (m_members[r].status != ML_OK)
            {
                char buf[512];
                m_members[r].scene->GetError(buf, sizeof(buf));
                throw std::runtime_error("device " + std::to_string(m_members[r].device->device) + ": " + buf);
            }
        if (!gather_error.empty()) throw std::runtime_error(gather_error);
        if (*dev0.wait_timeouts != before)
            throw std::runtime_error("multi-GPU frame: a completion flag did not arrive within the timeout");
        srt_render_stats total{};
        for (uint32_t r = 0; r < n; r++)
        {
            srt_render_stats s{};
            if (m_members[r].scene->GetStats(&s) != ML_OK)     // synchronises that device's stream
            {
                char buf[512];
                m_members[r].scene->GetError(buf, sizeof(buf));
                throw std::runtime_error(buf);
            }
            total.primary_rays += s.primary_rays;
            total.shadow_rays += s.shadow_rays;
            total.box_tests += s.box_tests;
            total.tri_tests += s.tri_tests;
            total.grid_build_ms = std::max(total.grid_build_ms, s.grid_build_ms);
            total.light_grid_build_ms = std::max(total.light_grid_build_ms, s.light_grid_build_ms);
        }
        dev0.MakeCurrent();
        ML_CUDA_CHECK(cudaEventElapsedTime(&total.kernel_ms, dev0.ev_start, m_ev_end));
        if (stats) *stats = total;
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        m_error = e.what();
        return ML_FAIL;
    }
}

} // namespace ML

extern "C" {

ML_API_ENTRY srt_scene_group srtCreateSceneGroup(int const* devices, uint32_t n_devices, size_t n_triangles,
                                                 float const* vertices, float const* albedo, int build_bvh,
                                                 char* error, size_t error_size)
{
    try
    {
        ML::FillBuffer(error, error_size, "");
        return ML::SceneGroup::MakeHandle(
            new ML::SceneGroup(devices, n_devices, n_triangles, vertices, albedo, build_bvh != 0));
    }
    catch (std::exception const& e)
    {
        ML::FillBuffer(error, error_size, e.what());
        return nullptr;
    }
}

ML_API_ENTRY ml_status srtRenderGroup(srt_scene_group group, srt_camera const* camera, srt_light const* light,
                                      srt_render_params const* params, float* rgb_host, srt_render_stats* stats)
{
    if (ML::SceneGroup::FromHandle(group) == nullptr) return ML_FAIL;
    return ML::SceneGroup::FromHandle(group)->Render(camera, light, params, rgb_host, stats);
}

ML_API_ENTRY void* srtSceneGroupFrame(srt_scene_group group)
{
    return ML::SceneGroup::FromHandle(group) ? ML::SceneGroup::FromHandle(group)->Frame() : nullptr;
}

ML_API_ENTRY uint32_t srtSceneGroupSize(srt_scene_group group)
{
    return ML::SceneGroup::FromHandle(group) ? ML::SceneGroup::FromHandle(group)->Size() : 0u;
}

ML_API_ENTRY char* srtGetSceneGroupError(srt_scene_group group, char* buffer, size_t buffer_size)
{
    if (ML::SceneGroup::FromHandle(group) == nullptr)
        return ML::FillBuffer(buffer, buffer_size, "Bad scene group handle");
    return ML::SceneGroup::FromHandle(group)->GetError(buffer, buffer_size);
}

ML_API_ENTRY void srtReleaseSceneGroup(srt_scene_group group)
{
    delete ML::SceneGroup::FromHandle(group);
}

} // extern "C"
