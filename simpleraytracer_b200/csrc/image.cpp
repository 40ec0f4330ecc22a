// image.cpp -- image object and its five C entry points.
// Contract: reference image.cpp:22-75 (validation messages, zero fill, Map
// returns a stable pointer + byte size, Unmap checks pointer identity) and
// :80-113 (C wrappers).  New here: pinned host memory + device twin with
// validity flags, so frames rendered or inferred on the GPU stay in HBM until
// somebody maps them.
#include "image.h"

#include "context.h"
#include "dtype.h"
#include "model_runner_ext.h"

namespace ML {

Image::Image(DeviceStatePtr device, ml_image_info const* info) : m_device(std::move(device))
{
    if (info == nullptr) throw std::runtime_error("Bad image information argument");

    size_t const item_size = DataTypeSize(info->dtype);

    ForEachDim([info](auto dim, char const* name) {
        if (info->*dim == 0)
            throw std::runtime_error(std::string("Unspecified image ") + name + " dimension");
        return true;
    });

    m_info = *info;
    m_bytes = m_info.width * m_info.height * m_info.channels * item_size;

    m_device->MakeCurrent();
    ML_CUDA_CHECK(cudaMallocHost(&m_host, m_bytes));
    cudaError_t err = cudaMalloc(&m_dev, m_bytes);
    if (err != cudaSuccess)
    {
        cudaFreeHost(m_host);
        m_host = nullptr;
        throw std::runtime_error(std::string("CUDA error: ") + cudaGetErrorString(err)
                                 + " allocating " + std::to_string(m_bytes) + " bytes of device memory");
    }
    std::memset(m_host, 0, m_bytes);
    ML_CUDA_CHECK(cudaMemsetAsync(m_dev, 0, m_bytes, m_device->stream));
    ML_CUDA_CHECK(cudaStreamSynchronize(m_device->stream));
    m_valid = kHostValid | kDeviceValid;
}

Image::~Image()
{
    cudaSetDevice(m_device->device);
    if (m_dev) cudaFree(m_dev);
    if (m_host) cudaFreeHost(m_host);
}

ml_status Image::GetInfo(ml_image_info* info) const
{
    if (info == nullptr) return ML_FAIL;
    *info = m_info;
    return ML_OK;
}

void* Image::Map(size_t* size)
{
    if (!(m_valid & kHostValid))
    {
        m_device->MakeCurrent();
        ML_CUDA_CHECK(cudaMemcpyAsync(m_host, m_dev, m_bytes, cudaMemcpyDeviceToHost, m_device->stream));
        ML_CUDA_CHECK(cudaStreamSynchronize(m_device->stream));
        m_valid |= kHostValid;
    }
    m_host_escaped = true;
    if (size != nullptr) *size = m_bytes;
    return m_host;
}

ml_status Image::Unmap(void* data)
{
    if (data != m_host) return ML_FAIL;
    // The caller may have written through the pointer: host copy wins.
    m_valid = kHostValid;
    return ML_OK;
}

void* Image::DeviceRead(cudaStream_t stream)
{
    if (m_valid & kHostValid)
    {
        ML_CUDA_CHECK(cudaMemcpyAsync(m_dev, m_host, m_bytes, cudaMemcpyHostToDevice, stream));
        m_valid |= kDeviceValid;
    }
    return m_dev;
}

void* Image::DeviceWrite()
{
    return m_dev;
}

void Image::CommitDeviceWrite(cudaStream_t stream)
{
    m_valid = kDeviceValid;
    if (m_host_escaped)
    {
        ML_CUDA_CHECK(cudaMemcpyAsync(m_host, m_dev, m_bytes, cudaMemcpyDeviceToHost, stream));
        m_valid |= kHostValid; // valid once the caller has synchronised `stream`
    }
}

} // namespace ML

extern "C" {

ML_API_ENTRY ml_status mlGetImageInfo(ml_image image, ml_image_info* info)
{
    if (ML::Image::FromHandle(image) == nullptr) return ML_FAIL;
    return ML::Image::FromHandle(image)->GetInfo(info);
}

ML_API_ENTRY void* mlMapImage(ml_image image, size_t* size)
{
    if (ML::Image::FromHandle(image) == nullptr) return nullptr;
    try
    {
        return ML::Image::FromHandle(image)->Map(size);
    }
    catch (std::exception const& e)
    {
        std::fprintf(stderr, "mlMapImage: %s\n", e.what());
        return nullptr;
    }
}

ML_API_ENTRY ml_status mlUnmapImage(ml_image image, void* data)
{
    if (ML::Image::FromHandle(image) == nullptr) return ML_FAIL;
    return ML::Image::FromHandle(image)->Unmap(data);
}

ML_API_ENTRY void mlReleaseImage(ml_image image)
{
    delete ML::Image::FromHandle(image);
}

ML_API_ENTRY ml_status mlxIpcExportImage(ml_image image, void* handle_out)
{
    ML::Image* img = ML::Image::FromHandle(image);
    if (img == nullptr || handle_out == nullptr) return ML_FAIL;
    static_assert(sizeof(cudaIpcMemHandle_t) == MLX_IPC_HANDLE_BYTES, "IPC handle size");
    try
    {
        img->Device()->MakeCurrent();
        cudaIpcMemHandle_t h;
        ML_CUDA_CHECK(cudaIpcGetMemHandle(&h, img->DeviceWrite()));
        std::memcpy(handle_out, &h, sizeof(h));
        img->MarkDeviceAuthoritative();
        return ML_OK;
    }
    catch (std::exception const& e)
    {
        std::fprintf(stderr, "mlxIpcExportImage: %s\n", e.what());
        return ML_FAIL;
    }
}

ML_API_ENTRY void* mlxIpcOpen(ml_context context, void const* handle)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr || handle == nullptr) return nullptr;
    ctx->ClearError();
    try
    {
        ctx->Device()->MakeCurrent();
        cudaIpcMemHandle_t h;
        std::memcpy(&h, handle, sizeof(h));
        void* p = nullptr;
        ML_CUDA_CHECK(cudaIpcOpenMemHandle(&p, h, cudaIpcMemLazyEnablePeerAccess));
        return p;
    }
    catch (std::exception const& e)
    {
        ctx->SetError(e.what());
        return nullptr;
    }
}

ML_API_ENTRY void mlxIpcClose(ml_context context, void* device_pointer)
{
    ML::Context* ctx = ML::Context::FromHandle(context);
    if (ctx == nullptr || device_pointer == nullptr) return;
    cudaSetDevice(ctx->Device()->device);
    cudaIpcCloseMemHandle(device_pointer);
}

ML_API_ENTRY void* mlxImageDevicePointer(ml_image image, int for_write, void* stream)
{
    ML::Image* img = ML::Image::FromHandle(image);
    if (img == nullptr) return nullptr;
    try
    {
        img->Device()->MakeCurrent();
        cudaStream_t s = stream ? static_cast<cudaStream_t>(stream) : img->Device()->stream;
        if (for_write)
        {
            void* p = img->DeviceWrite();
            img->MarkDeviceAuthoritative();
            return p;
        }
        return img->DeviceRead(s);
    }
    catch (std::exception const& e)
    {
        std::fprintf(stderr, "mlxImageDevicePointer: %s\n", e.what());
        return nullptr;
    }
}

} // extern "C"
