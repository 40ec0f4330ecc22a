// nn_kernels.h -- launchers of the hand-written inference kernels that stand
// where tensorflow::Session::Run stood (reference model.cpp:339).  Tensors are
// dense NHWC with N = 1 (the reference's input is {1,H,W,C}, model.cpp:208-213),
// convolution filters are HWIO as in TensorFlow's Conv2D.
#pragma once

#include <cuda_runtime.h>

#include <cstddef>
#include <cstdint>
#include <vector>

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
    float const* residual; // fp32 [OH*OW*Cout] added AFTER the activation (a fused Add), or nullptr
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
// ---- the ops around the convolutions of an encoder / decoder graph (TensorFlow op definitions, NHWC, N = 1)
// out = act(in * scale[c] + bias[c]); scale / bias may be nullptr.  In place allowed.
void ScaleBiasAct(float const* in, float const* scale, float const* bias, float* out, size_t n, int C, Act act,
                  float alpha, cudaStream_t stream);
// mode 0: a + b, 1: a - b, 2: a * b, then the activation.  In place allowed.
void BinaryAct(float const* a, float const* b, float* out, size_t n, int mode, Act act, float alpha, cudaStream_t stream);
struct PoolDesc { int H, W, C, OH, OW, K, stride, pad_top, pad_left, is_max; };
void Pool2d(float const* in, float* out, PoolDesc const& d, cudaStream_t stream);
struct ResizeDesc { int H, W, C, OH, OW, bilinear, align_corners, half_pixel; float scale_y, scale_x; };
void Resize(float const* in, float* out, ResizeDesc const& d, cudaStream_t stream);
void DepthToSpace(float const* in, float* out, int H, int W, int C, int block, cudaStream_t stream);
// mode 0: CONSTANT (zeros), 1: REFLECT, 2: SYMMETRIC
void Pad2d(float const* in, float* out, int H, int W, int C, int top, int left, int OH, int OW, int mode,
           cudaStream_t stream);
void HalfToFloat(void const* in, float* out, size_t n, cudaStream_t stream);
// fp32 direct convolution writing bf16 NHWC (feeds the tensor-core layers).
void Conv2dFp32ToBf16(float const* in, float const* w, float const* bias, void* out_bf16,
                      ConvDesc const& d, cudaStream_t stream);

// ---- bf16 tensor-core path (conv_tc_kernels.cu): 3x3, stride 1, SAME, tcgen05 + TMEM + TMA.
// Activations bf16 NHWC, filter repacked [tap][Cout padded to 16][Cin] bf16, fp32 accumulation.
struct ConvTc; // one planned layer: tensor maps bound to fixed device buffers
// split = the bf16x3 (fp32-grade) variant: operands stored as [hi | lo] bf16 halves, three products per
// K chunk (hi*hi + lo*hi + hi*lo) into the fp32 accumulator, fp32 output.
bool ConvTcSupported(int KH, int KW, int stride, int dilation, bool same_padding, int Cin, int Cout,
                     bool split = false);
int ConvTcPaddedCout(int Cout);
// Cin_pad >= Cin: input channels beyond Cin get zero filter taps (3-channel first layer padded to 16).
std::vector<uint16_t> ConvTcPackWeights(float const* hwio, int Cin, int Cout, int Cin_pad);
// Split filter: [tap][Cout padded][hi: Cin_pad | lo: Cin_pad] with w = hi + lo.
std::vector<uint16_t> ConvTcPackWeightsSplit(float const* hwio, int Cin, int Cout, int Cin_pad);
// fp32 [pixels][C] -> bf16 [pixels][hi: Cp | lo: Cp], x = hi + lo to 16 mantissa bits, channels >= C zero.
// Cp is a multiple of 8, Cp >= C.
void SplitToBf16HiLo(float const* in, void* out_bf16, size_t pixels, int C, int Cp, cudaStream_t stream);
// fp32 [pixels][C <= 16] -> bf16 [pixels][16], zero padded (input of a padded first layer).
void PadToBf16x16(float const* in, void* out_bf16, size_t pixels, int C, cudaStream_t stream);
// out_bf16: bf16 NHWC output (Cout % 16 == 0) for the next tensor-core layer; else fp32 NHWC.
// d_out_bf16: the (fixed) bf16 output buffer the TMA store is bound to; ignored for fp32 output,
// whose pointer is given to ConvTcRun.
ConvTc* ConvTcCreate(void const* d_in_bf16, void const* d_w_packed, void* d_out_bf16, int H, int W, int Cin,
                     int Cout, bool out_bf16, Act act, float alpha, int sm_count, bool split = false,
                     bool out_split = false);   // out_split: the epilogue writes bf16 [hi | lo] halves (Cout % 16 == 0)
void ConvTcRun(ConvTc const* op, float const* bias, void* out, cudaStream_t stream);
void ConvTcDestroy(ConvTc* op);
void FloatToHalf(float const* in, void* out, size_t n, cudaStream_t stream);

} // namespace nn
} // namespace ML
