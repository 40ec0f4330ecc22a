// context.cpp -- context object and its five C entry points.
// Behavioural contract: reference context.cpp:20-54 (factories clear the
// context error, convert any exception into the error string and return NULL)
// and :59-105 (C wrappers).  New here: the context selects the GPU
// (MLB200_DEVICE, else the calling thread's current device) and owns the
// stream and events every child object uses.
#include "context.h"

#include "image.h"
#include "model.h"
#include "model_runner_ext.h"

#include <algorithm>
#include <condition_variable>
#include <mutex>
#include <thread>
#include <vector>

namespace ML {

DeviceState::DeviceState(int requested_device)
{
    int count = 0;
    cudaError_t err = cudaGetDeviceCount(&count);
    if (err != cudaSuccess || count == 0)
    {
        throw std::runtime_error(std::string("No CUDA device available: ")
                                 + (err != cudaSuccess ? cudaGetErrorString(err) : "device count is 0")
                                 + " (this library has no CPU fallback)");
    }
    if (requested_device >= 0)
    {
        device = requested_device;
        if (device >= count)
            throw std::runtime_error("CUDA device ordinal out of range: " + std::to_string(device));
    }
    else if (char const* env = std::getenv("MLB200_DEVICE"))
    {
        device = std::atoi(env);
        if (device < 0 || device >= count)
            throw std::runtime_error("MLB200_DEVICE out of range: " + std::string(env));
    }
    else
    {
        ML_CUDA_CHECK(cudaGetDevice(&device));
    }
    ML_CUDA_CHECK(cudaSetDevice(device));
    ML_CUDA_CHECK(cudaGetDeviceProperties(&prop, device));
    ML_CUDA_CHECK(cudaStreamCreateWithFlags(&stream, cudaStreamNonBlocking));
    ML_CUDA_CHECK(cudaStreamCreateWithFlags(&copy_in, cudaStreamNonBlocking));
    ML_CUDA_CHECK(cudaStreamCreateWithFlags(&copy_out, cudaStreamNonBlocking));
    ML_CUDA_CHECK(cudaEventCreate(&ev_start));
    ML_CUDA_CHECK(cudaEventCreate(&ev_stop));
    ML_CUDA_CHECK(cudaHostAlloc(reinterpret_cast<void**>(&wait_timeouts), sizeof(uint32_t), cudaHostAllocMapped));
    *wait_timeouts = 0;
}

DeviceState::~DeviceState()
{
    // Destructors must not throw; errors here have nowhere to go.
    cudaSetDevice(device);
    if (ev_start) cudaEventDestroy(ev_start);
    if (ev_stop) cudaEventDestroy(ev_stop);
    if (copy_in) cudaStreamDestroy(copy_in);
    if (copy_out) cudaStreamDestroy(copy_out);
    if (stream) cudaStreamDestroy(stream);
    if (wait_timeouts) cudaFreeHost(wait_timeouts);
    if (gather_stage) cudaFreeHost(gather_stage);
    for (cudaEvent_t e : gather_events)
        if (e) cudaEventDestroy(e);
}

namespace {
// Fixed set of workers; one job at a time (callers are serialised by a mutex: frames are copied one after another).
class CopyPool
{
public:
    static CopyPool& Instance()
    {
        static CopyPool* pool = new CopyPool;     // leaked on purpose: workers may outlive static destruction order
        return *pool;
    }
    void Copy(char* dst, char const* src, size_t bytes)
    {
        size_t const n = m_threads.size() + 1;    // the caller copies a slice too
        std::lock_guard<std::mutex> serial(m_serial);
        size_t const slice = ((bytes + n - 1) / n + 4095) / 4096 * 4096;
        {
            std::lock_guard<std::mutex> lock(m_mutex);
            m_dst = dst; m_src = src; m_bytes = bytes; m_slice = slice;
            m_pending = m_threads.size();
            m_job++;
        }
        m_wake.notify_all();
        CopySlice(0);
        std::unique_lock<std::mutex> lock(m_mutex);
        m_done.wait(lock, [&] { return m_pending == 0; });
    }

private:
    CopyPool()
    {
        unsigned want = std::thread::hardware_concurrency() / 2;
        if (char const* e = std::getenv("SRT_HOST_COPY_THREADS")) want = static_cast<unsigned>(std::atoi(e));
        want = std::max(1u, std::min(want, 8u));
        for (unsigned t = 1; t < want; t++) m_threads.emplace_back([this, t] { Worker(t); });
        for (auto& th : m_threads) th.detach();
    }
    void CopySlice(size_t index)
    {
        size_t const begin = index * m_slice;
        if (begin < m_bytes) std::memcpy(m_dst + begin, m_src + begin, std::min(m_slice, m_bytes - begin));
    }
    void Worker(size_t index)
    {
        uint64_t seen = 0;
        for (;;)
        {
            {
                std::unique_lock<std::mutex> lock(m_mutex);
                m_wake.wait(lock, [&] { return m_job != seen; });
                seen = m_job;
            }
            CopySlice(index);
            {
                std::lock_guard<std::mutex> lock(m_mutex);
                m_pending--;
            }
            m_done.notify_one();
        }
    }
    std::vector<std::thread> m_threads;
    std::mutex m_mutex, m_serial;
    std::condition_variable m_wake, m_done;
    char* m_dst = nullptr;
    char const* m_src = nullptr;
    size_t m_bytes = 0, m_slice = 0, m_pending = 0;
    uint64_t m_job = 0;
};
} // namespace

void ParallelMemcpy(void* dst, void const* src, size_t bytes)
{
    if (bytes < (size_t(1) << 20))     // below ~1 MiB one thread is done before the pool has woken up
    {
        std::memcpy(dst, src, bytes);
        return;
    }
    CopyPool::Instance().Copy(static_cast<char*>(dst), static_cast<char const*>(src), bytes);
}

Context::Context(int device) : m_device(std::make_shared<DeviceState>(device)) {}

ml_image Context::CreateImage(ml_image_info const* info)
{
    m_error.clear();
    try
    {
        return Image::MakeHandle(new Image(m_device, info));
    }
    catch (std::exception const& e)
    {
        m_error = e.what();
        return ML_INVALID_HANDLE;
    }
}

ml_model Context::CreateModel(ml_model_params const* params)
{
    m_error.clear();
    try
    {
        return Model::MakeHandle(new Model(m_device, params));
    }
    catch (std::exception const& e)
    {
        m_error = e.what();
        return ML_INVALID_HANDLE;
    }
}

char* Context::GetError(char* buffer, size_t buffer_size) const
{
    return FillBuffer(buffer, buffer_size, m_error);
}

} // namespace ML

extern "C" {

ML_API_ENTRY ml_context mlCreateContext()
{
    try
    {
        return ML::Context::MakeHandle(new ML::Context);
    }
    catch (std::exception const& e)
    {
        // The C API has no channel for this message (the reference returns a
        // bare NULL, context.cpp:59-69); say why on stderr so that a missing
        // GPU is never mistaken for success.
        std::fprintf(stderr, "mlCreateContext: %s\n", e.what());
        return ML_INVALID_HANDLE;
    }
    catch (...)
    {
        return ML_INVALID_HANDLE;
    }
}

ML_API_ENTRY ml_context mlxCreateContextOnDevice(int device)
{
    try
    {
        if (device < 0) throw std::runtime_error("negative device ordinal");
        return ML::Context::MakeHandle(new ML::Context(device));
    }
    catch (std::exception const& e)
    {
        std::fprintf(stderr, "mlxCreateContextOnDevice: %s\n", e.what());
        return ML_INVALID_HANDLE;
    }
}

ML_API_ENTRY char* mlGetContextError(ml_context context, char* buffer, size_t buffer_size)
{
    if (ML::Context::FromHandle(context) == nullptr)
        return ML::FillBuffer(buffer, buffer_size, "Bad context handle");
    return ML::Context::FromHandle(context)->GetError(buffer, buffer_size);
}

ML_API_ENTRY void mlReleaseContext(ml_context context)
{
    delete ML::Context::FromHandle(context);
}

ML_API_ENTRY ml_image mlCreateImage(ml_context context, ml_image_info const* info)
{
    if (ML::Context::FromHandle(context) == nullptr) return ML_INVALID_HANDLE;
    return ML::Context::FromHandle(context)->CreateImage(info);
}

ML_API_ENTRY ml_model mlCreateModel(ml_context context, ml_model_params const* params)
{
    if (ML::Context::FromHandle(context) == nullptr) return ML_INVALID_HANDLE;
    return ML::Context::FromHandle(context)->CreateModel(params);
}

} // extern "C"
