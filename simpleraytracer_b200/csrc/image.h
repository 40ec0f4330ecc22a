// image.h -- ML::Image: dense HWC buffer (reference image.h:10-25) kept as a
// pinned host allocation plus a device-resident twin in HBM.
#pragma once

#include "common.h"

namespace ML {

class Image
{
public:
    static ml_image MakeHandle(Image* i) { return reinterpret_cast<ml_image>(i); }
    static Image* FromHandle(ml_image i) { return reinterpret_cast<Image*>(i); }

    Image(DeviceStatePtr device, ml_image_info const* info);
    ~Image();
    Image(Image const&) = delete;
    Image& operator=(Image const&) = delete;

    ml_status GetInfo(ml_image_info* info) const;
    void* Map(size_t* size);
    ml_status Unmap(void* data);

    // ---- device side (used by Model::Infer and the renderer) -------------
    ml_image_info const& Info() const { return m_info; }
    size_t Bytes() const { return m_bytes; }
    DeviceStatePtr const& Device() const { return m_device; }

    // Device pointer holding the current contents: uploads the host copy first
    // whenever the host copy is valid (the host is authoritative while valid,
    // because a caller may legally keep writing through a mapped pointer).
    void* DeviceRead(cudaStream_t stream);
    // Device pointer that is about to be overwritten entirely on `stream`.
    void* DeviceWrite();
    // After a device write has been enqueued: the device copy is the only
    // valid one.  If the host pointer has ever been handed out, the data is
    // copied back eagerly (asynchronously on `stream`; the caller syncs).
    void CommitDeviceWrite(cudaStream_t stream);
    // An external writer owns the device copy now (mlxImageDevicePointer): the
    // next Map() copies it back.
    void MarkDeviceAuthoritative() { m_valid = kDeviceValid; }
    bool HostPointerEscaped() const { return m_host_escaped; }

private:
    enum : unsigned { kHostValid = 1u, kDeviceValid = 2u };

    DeviceStatePtr m_device;
    ml_image_info m_info{};
    size_t m_bytes = 0;
    void* m_host = nullptr;   // cudaMallocHost (pinned)
    void* m_dev = nullptr;    // cudaMalloc
    unsigned m_valid = 0;
    bool m_host_escaped = false;
};

} // namespace ML
