// nn_kernels.h -- launchers of the hand-written inference kernels that stand
// where tensorflow::Session::Run stood (reference model.cpp:339).  Tensors are
// dense NHWC with N = 1 (the reference's input is {1,H,W,C}, model.cpp:208-213),
// convolution filters are HWIO as in TensorFlow's Conv2D.
#pragma once

#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>

namespace ML {
namespace nn {

enum class Act : int { None = 0, Relu, LeakyRelu, Relu6, Sigmoid, Tanh };

struct ConvDesc
{
    int H, W, Cin;      // input
    int OH, OW, Cout;   // output
    int KH, KW;
    int stride;         // same in y and x
    int dilation;       // same in y and x
    int pad_top, pad_left;
    Act act;
    float alpha;        // LeakyRelu slope
};

// fp32 direct convolution + bias + activation, one launch.
// in: H*W*Cin, w: KH*KW*Cin*Cout (HWIO), bias: Cout or nullptr, out: OH*OW*Cout.
void Conv2dFp32(float const* in, float const* w, float const* bias, float* out,
                ConvDesc const& d, cudaStream_t stream);

// out[i] = act(in[i] + bias[i % C]) ; bias may be nullptr.
void BiasAct(float const* in, float const* bias, float* out, size_t n, int C, Act act, float alpha,
             cudaStream_t stream);
// out[i] = act(a[i] + b[i])
void AddAct(float const* a, float const* b, float* out, size_t n, Act act, float alpha,
            cudaStream_t stream);
// Copies a [pixels, C] tensor into channels [c0, c0+C) of a [pixels, Ctotal] tensor.
void ConcatChannels(float const* in, float* out, size_t pixels, int C, int Ctotal, int c0,
                    cudaStream_t stream);
void HalfToFloat(void const* in, float* out, size_t n, cudaStream_t stream);
void FloatToHalf(float const* in, void* out, size_t n, cudaStream_t stream);

} // namespace nn
} // namespace ML
