// common.h -- shared host-side helpers of the B200 model runner / renderer.
#pragma once

#include "model_runner.h"

#include <cuda_runtime.h>

#include <cstdlib>
#include <cstring>
#include <memory>
#include <sstream>
#include <stdexcept>
#include <string>

namespace ML {

#define ML_CUDA_CHECK(expr)                                                        \
    do {                                                                           \
        cudaError_t err__ = (expr);                                                \
        if (err__ != cudaSuccess) {                                                \
            std::ostringstream os__;                                               \
            os__ << "CUDA error: " << cudaGetErrorString(err__) << " (" #expr ") " \
                 << __FILE__ << ":" << __LINE__;                                   \
            throw std::runtime_error(os__.str());                                  \
        }                                                                          \
    } while (0)

// Visits width, height, channels in that order and stops at the first visitor
// that returns false (same order/short-circuit as reference utils.h:12-18).
template <class Visitor>
bool ForEachDim(Visitor const& visitor)
{
    using DimMember = size_t ml_image_info::*;
    struct Dim { DimMember member; char const* name; };
    static constexpr Dim dims[] = {
        {&ml_image_info::width, "width"},
        {&ml_image_info::height, "height"},
        {&ml_image_info::channels, "channels"},
    };
    for (auto const& d : dims)
        if (!visitor(d.member, d.name)) return false;
    return true;
}

// Error-string export used by ml*GetError.  Intentional divergence from the
// reference's FillBuffer (utils.h:20-28): the full message
// is written, truncated to buffer_size-1 characters plus NUL; an empty message
// yields ""; buffer_size == 0 writes nothing; never throws.
inline char* FillBuffer(char* buffer, size_t buffer_size, std::string const& message) noexcept
{
    if (buffer == nullptr || buffer_size == 0) return buffer;
    size_t n = message.size() < buffer_size - 1 ? message.size() : buffer_size - 1;
    std::memcpy(buffer, message.data(), n);
    buffer[n] = '\0';
    return buffer;
}

// The CUDA side of a context: device ordinal, one compute stream, two timing
// events.  Shared (ref-counted) by the context and every image/model/scene it
// creates, because the reference lets the caller release the context before
// its children (context.cpp:102-105 holds no references).
struct DeviceState
{
    int device = 0;
    cudaStream_t stream = nullptr;
    cudaStream_t copy_in = nullptr;   // H2D of the next image while the current one computes (mlxInferBatch)
    cudaStream_t copy_out = nullptr;  // D2H of the previous result
    cudaEvent_t ev_start = nullptr;
    cudaEvent_t ev_stop = nullptr;
    cudaDeviceProp prop{};
    // mapped pinned word: number of device-side flag waits that gave up (mlxStreamWaitFlags timeout)
    uint32_t* wait_timeouts = nullptr;
    // pinned staging + events of srtGatherFrameHost (pageable destination frames), grown on demand
    void* gather_stage = nullptr;
    size_t gather_stage_bytes = 0;
    cudaEvent_t gather_events[16] = {};

    explicit DeviceState(int requested_device = -1); // -1: MLB200_DEVICE, else current device
    ~DeviceState();
    DeviceState(DeviceState const&) = delete;
    DeviceState& operator=(DeviceState const&) = delete;

    void MakeCurrent() const { ML_CUDA_CHECK(cudaSetDevice(device)); }
    void Sync() const { ML_CUDA_CHECK(cudaStreamSynchronize(stream)); }
};

using DeviceStatePtr = std::shared_ptr<DeviceState>;

// memcpy of a large piece (pinned staging -> the caller's pageable frame) split over a small PERSISTENT pool of host
// threads (created on first use, alive until exit: starting threads per call cost more than it saved, r01).  Pieces
// below 2 MiB are copied by the calling thread.  A 99.5 MB 4K frame is host-copy bound on one thread (~9 GB/s).
void ParallelMemcpy(void* dst, void const* src, size_t bytes);

} // namespace ML
