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
