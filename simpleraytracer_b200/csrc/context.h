// context.h -- ML::Context: factory + last-error string (reference
// context.h:13-25) that additionally owns the CUDA device state.
#pragma once

#include "common.h"

namespace ML {

class Context
{
public:
    explicit Context(int device = -1);

    static ml_context MakeHandle(Context* c) { return reinterpret_cast<ml_context>(c); }
    static Context* FromHandle(ml_context c) { return reinterpret_cast<Context*>(c); }

    ml_image CreateImage(ml_image_info const* info);
    ml_model CreateModel(ml_model_params const* params);
    char* GetError(char* buffer, size_t buffer_size) const;

    DeviceStatePtr const& Device() const { return m_device; }
    void ClearError() { m_error.clear(); }
    void SetError(std::string const& message) { m_error = message; }

private:
    DeviceStatePtr m_device;
    std::string m_error;
};

} // namespace ML
