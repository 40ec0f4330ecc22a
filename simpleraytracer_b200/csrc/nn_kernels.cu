// nn_kernels.cu -- fp32 inference kernels (CUDA cores) for sm_100a.
//
// conv2d_tile_kernel: stride-1 direct convolution.  A block owns a TH x TW
// output tile and up to 4*CG output channels; the input tile (with halo) and
// the filter slice for CI_T input channels are staged in shared memory, every
// thread keeps PXT pixels x 4 channels of accumulators in registers, reads the
// filter as one LDS.128 per (kh,kw,ci) and slides a register window along the
// row so each input value is loaded once per kh.  Bias and activation are
// fused into the store.  Bound: FP32 FMA pipe (2*KH*KW*Cin*Cout FLOP/pixel).
//
// conv2d_naive_kernel: any stride/dilation, one thread per output value.
#include "nn_kernels.h"

#include "common.h"

#include <cuda_bf16.h>
#include <cuda_fp16.h>

namespace ML {
namespace nn {
namespace {

__device__ __forceinline__ float apply_act(float v, Act act, float alpha)
{
    switch (act)
    {
        case Act::Relu: return v > 0.f ? v : 0.f;
        case Act::LeakyRelu: return v > 0.f ? v : v * alpha;
        case Act::Relu6: return v < 0.f ? 0.f : (v > 6.f ? 6.f : v);
        case Act::Sigmoid: return 1.0f / (1.0f + expf(-v));
        case Act::Tanh: return tanhf(v);
        default: return v;
    }
}

// Packed pair of fp32 in one 64-bit register: sm_100 fma.rn.f32x2 (SASS FFMA2) does two FMAs per
// issue slot, which leaves issue slots for the shared-memory loads of a direct convolution.
typedef unsigned long long F2;
__device__ __forceinline__ F2 pack2(float a, float b)
{
    F2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ float2 unpack2(F2 v)
{
    float2 r;
    asm("mov.b64 {%0, %1}, %2;" : "=f"(r.x), "=f"(r.y) : "l"(v));
    return r;
}
__device__ __forceinline__ F2 fma2(F2 a, F2 b, F2 c)
{
    F2 d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

template <int KS, int TH, int TW, int PXT, int CG, int CI_T, typename OutT, int CPT = 4>
__global__ void __launch_bounds__(256, 2)
conv2d_tile_kernel(float const* __restrict__ in, float const* __restrict__ w,
                   float const* __restrict__ bias, OutT* __restrict__ out, ConvDesc d)
{
    constexpr int IH = TH + KS - 1, IW = TW + KS - 1;
    constexpr int CO_T = CPT * CG;   // CPT = 4 or 8 output channels per thread
    constexpr int GROUPS_PER_ROW = TW / PXT;
    static_assert(CPT == 4 || CPT == 8, "channels per thread");
    static_assert((TH * TW / PXT) * CG == 256, "thread mapping must cover 256 threads");

    __shared__ float s_in[IH][IW][CI_T + 1]; // +1: pixel groups of a warp land in different banks
    __shared__ __align__(16) float s_w[KS * KS][CI_T][CO_T];

    int const tid = threadIdx.x;
    int const cg = tid % CG;
    int const pg = tid / CG;
    int const row = pg / GROUPS_PER_ROW;
    int const col0 = (pg % GROUPS_PER_ROW) * PXT;
    int const ty0 = blockIdx.y * TH, tx0 = blockIdx.x * TW;
    int const cot0 = blockIdx.z * CO_T;

    F2 acc2[PXT][CPT / 2]; // channel pairs (0,1), (2,3), ... of every pixel
#pragma unroll
    for (int i = 0; i < PXT; i++)
#pragma unroll
        for (int j = 0; j < CPT / 2; j++) acc2[i][j] = pack2(0.f, 0.f);

    // Staging is software-pipelined through registers: the next channel slice is fetched from global
    // memory (float4 wherever the layout allows) BEFORE the FMAs of the current slice, and written to
    // shared memory after them, so the global-load latency hides behind ~2300 FFMA2 per thread.
    constexpr int VI = CI_T % 4 == 0 ? 4 : 1;                  // input channels per fetch
    constexpr int VW = CO_T % 4 == 0 ? 4 : 1;                  // filter output channels per fetch
    constexpr int IN_ITEMS = IH * IW * (CI_T / VI), W_ITEMS = KS * KS * CI_T * (CO_T / VW);
    constexpr int IN_PER = (IN_ITEMS + 255) / 256, W_PER = (W_ITEMS + 255) / 256;
    bool const vec_in = VI == 4 && d.Cin % 4 == 0;            // 16-byte aligned pixels, no ragged channel tail
    bool const vec_w = VW == 4 && d.Cout % 4 == 0;
    float4 rin[IN_PER], rw[W_PER];
    auto fetch = [&](int c0) {
#pragma unroll
        for (int k = 0; k < IN_PER; k++)
        {
            int const idx = tid + 256 * k;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < IN_ITEMS)
            {
                int const cv = idx % (CI_T / VI), px = idx / (CI_T / VI);
                int const r = px / IW, c = px % IW;
                int const y = ty0 + r - d.pad_top, x = tx0 + c - d.pad_left;
                if (y >= 0 && y < d.H && x >= 0 && x < d.W)
                {
                    float const* src = in + (static_cast<size_t>(y) * d.W + x) * d.Cin + c0 + cv * VI;
                    if (vec_in)
                    {
                        if (c0 + cv * VI < d.Cin) v = __ldg(reinterpret_cast<float4 const*>(src));
                    }
                    else
                    {
                        if (c0 + cv * VI + 0 < d.Cin) v.x = __ldg(src + 0);
                        if (VI == 4 && c0 + cv * VI + 1 < d.Cin) v.y = __ldg(src + 1);
                        if (VI == 4 && c0 + cv * VI + 2 < d.Cin) v.z = __ldg(src + 2);
                        if (VI == 4 && c0 + cv * VI + 3 < d.Cin) v.w = __ldg(src + 3);
                    }
                }
            }
            rin[k] = v;
        }
#pragma unroll
        for (int k = 0; k < W_PER; k++)
        {
            int const idx = tid + 256 * k;
            float4 v = make_float4(0.f, 0.f, 0.f, 0.f);
            if (idx < W_ITEMS)
            {
                int const cov = idx % (CO_T / VW), ci = (idx / (CO_T / VW)) % CI_T, kk = idx / ((CO_T / VW) * CI_T);
                int const co = cot0 + cov * VW;
                if (c0 + ci < d.Cin)
                {
                    float const* src = w + (static_cast<size_t>(kk) * d.Cin + c0 + ci) * d.Cout + co;
                    if (vec_w)
                    {
                        if (co < d.Cout) v = __ldg(reinterpret_cast<float4 const*>(src));
                    }
                    else
                    {
                        if (co + 0 < d.Cout) v.x = __ldg(src + 0);
                        if (VW == 4 && co + 1 < d.Cout) v.y = __ldg(src + 1);
                        if (VW == 4 && co + 2 < d.Cout) v.z = __ldg(src + 2);
                        if (VW == 4 && co + 3 < d.Cout) v.w = __ldg(src + 3);
                    }
                }
            }
            rw[k] = v;
        }
    };
    auto commit = [&]() {
#pragma unroll
        for (int k = 0; k < IN_PER; k++)
        {
            int const idx = tid + 256 * k;
            if (idx < IN_ITEMS)
            {
                int const cv = idx % (CI_T / VI), px = idx / (CI_T / VI);
                int const r = px / IW, c = px % IW;
                float* dst = &s_in[r][c][cv * VI];
                dst[0] = rin[k].x;
                if (VI == 4) { dst[1] = rin[k].y; dst[2] = rin[k].z; dst[3] = rin[k].w; }
            }
        }
#pragma unroll
        for (int k = 0; k < W_PER; k++)
        {
            int const idx = tid + 256 * k;
            if (idx < W_ITEMS)
            {
                int const cov = idx % (CO_T / VW), ci = (idx / (CO_T / VW)) % CI_T, kk = idx / ((CO_T / VW) * CI_T);
                if (VW == 4) *reinterpret_cast<float4*>(&s_w[kk][ci][cov * 4]) = rw[k];
                else s_w[kk][ci][cov] = rw[k].x;
            }
        }
    };

    fetch(0);
    for (int c0 = 0; c0 < d.Cin; c0 += CI_T)
    {
        commit();
        __syncthreads();
        if (c0 + CI_T < d.Cin) fetch(c0 + CI_T);

#pragma unroll
        for (int kh = 0; kh < KS; kh++)
        {
#pragma unroll 2
            for (int ci = 0; ci < CI_T; ci++)
            {
                F2 xin[PXT + KS - 1]; // (x, x)
#pragma unroll
                for (int j = 0; j < PXT + KS - 1; j++)
                {
                    float const x = s_in[row + kh][col0 + j][ci];
                    xin[j] = pack2(x, x);
                }
#pragma unroll
                for (int kw = 0; kw < KS; kw++)
                {
#pragma unroll
                    for (int q = 0; q < CPT / 4; q++)
                    {
                        float4 wv = *reinterpret_cast<float4 const*>(&s_w[kh * KS + kw][ci][cg * CPT + 4 * q]);
                        F2 const w01 = pack2(wv.x, wv.y), w23 = pack2(wv.z, wv.w);
#pragma unroll
                        for (int i = 0; i < PXT; i++)
                        {
                            acc2[i][2 * q + 0] = fma2(xin[i + kw], w01, acc2[i][2 * q + 0]);
                            acc2[i][2 * q + 1] = fma2(xin[i + kw], w23, acc2[i][2 * q + 1]);
                        }
                    }
                }
            }
        }
        __syncthreads();
    }

    int const oy = ty0 + row;
    if (oy >= d.OH) return;
#pragma unroll
    for (int q = 0; q < CPT / 4; q++)
    {
        int const co = cot0 + cg * CPT + 4 * q;
        float b[4];
#pragma unroll
        for (int j = 0; j < 4; j++) b[j] = (bias != nullptr && co + j < d.Cout) ? bias[co + j] : 0.f;
        bool const vec = (d.Cout % 4 == 0) && (co + 3 < d.Cout);
#pragma unroll
        for (int i = 0; i < PXT; i++)
        {
            float2 const a01 = unpack2(acc2[i][2 * q]), a23 = unpack2(acc2[i][2 * q + 1]);
            int ox = tx0 + col0 + i;
            if (ox >= d.OW) continue;
            OutT* o = out + (static_cast<size_t>(oy) * d.OW + ox) * d.Cout + co;
            float r0 = apply_act(a01.x + b[0], d.act, d.alpha);
            float r1 = apply_act(a01.y + b[1], d.act, d.alpha);
            float r2 = apply_act(a23.x + b[2], d.act, d.alpha);
            float r3 = apply_act(a23.y + b[3], d.act, d.alpha);
            if constexpr (sizeof(OutT) == 4)
            {
                if (d.residual != nullptr)
                {
                    // fused Add: same fp32 additions, in the same order, as the separate add kernel made
                    float const* res = d.residual + (static_cast<size_t>(oy) * d.OW + ox) * d.Cout + co;
                    if (vec)
                    {
                        float4 const rv = *reinterpret_cast<float4 const*>(res);
                        r0 += rv.x; r1 += rv.y; r2 += rv.z; r3 += rv.w;
                    }
                    else
                    {
                        if (co + 0 < d.Cout) r0 += res[0];
                        if (co + 1 < d.Cout) r1 += res[1];
                        if (co + 2 < d.Cout) r2 += res[2];
                        if (co + 3 < d.Cout) r3 += res[3];
                    }
                }
            }
            if (vec)
            {
                if constexpr (sizeof(OutT) == 4)
                {
                    *reinterpret_cast<float4*>(o) = make_float4(r0, r1, r2, r3);
                }
                else
                {
                    __nv_bfloat162 lo = __floats2bfloat162_rn(r0, r1), hi = __floats2bfloat162_rn(r2, r3);
                    *reinterpret_cast<uint2*>(o) = make_uint2(*reinterpret_cast<uint32_t*>(&lo),
                                                              *reinterpret_cast<uint32_t*>(&hi));
                }
            }
            else
            {
                if (co + 0 < d.Cout) o[0] = static_cast<OutT>(r0);
                if (co + 1 < d.Cout) o[1] = static_cast<OutT>(r1);
                if (co + 2 < d.Cout) o[2] = static_cast<OutT>(r2);
                if (co + 3 < d.Cout) o[3] = static_cast<OutT>(r3);
            }
        }
    }
}

__global__ void conv2d_naive_kernel(float const* __restrict__ in, float const* __restrict__ w,
                                    float const* __restrict__ bias, float* __restrict__ out, ConvDesc d)
{
    size_t const total = static_cast<size_t>(d.OH) * d.OW * d.Cout;
    size_t idx = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    if (idx >= total) return;
    int co = static_cast<int>(idx % d.Cout);
    size_t px = idx / d.Cout;
    int ox = static_cast<int>(px % d.OW), oy = static_cast<int>(px / d.OW);
    float acc = 0.f;
    for (int kh = 0; kh < d.KH; kh++)
    {
        int y = oy * d.stride + kh * d.dilation - d.pad_top;
        if (y < 0 || y >= d.H) continue;
        for (int kw = 0; kw < d.KW; kw++)
        {
            int x = ox * d.stride + kw * d.dilation - d.pad_left;
            if (x < 0 || x >= d.W) continue;
            float const* ip = in + (static_cast<size_t>(y) * d.W + x) * d.Cin;
            float const* wp = w + (static_cast<size_t>(kh) * d.KW + kw) * d.Cin * d.Cout + co;
            for (int ci = 0; ci < d.Cin; ci++) acc = fmaf(ip[ci], wp[static_cast<size_t>(ci) * d.Cout], acc);
        }
    }
    if (bias) acc += bias[co];
    acc = apply_act(acc, d.act, d.alpha);
    if (d.residual != nullptr) acc += d.residual[idx];
    out[idx] = acc;
}

template <int KS, int TH, int TW, int PXT, int CG, int CI_T, int CPT = 4, typename OutT>
void LaunchTile(float const* in, float const* w, float const* bias, OutT* out, ConvDesc const& d,
                cudaStream_t stream)
{
    dim3 grid((d.OW + TW - 1) / TW, (d.OH + TH - 1) / TH, (d.Cout + CPT * CG - 1) / (CPT * CG));
    conv2d_tile_kernel<KS, TH, TW, PXT, CG, CI_T, OutT, CPT><<<grid, 256, 0, stream>>>(in, w, bias, out, d);
}

template <int KS, int CI_T, typename OutT>
void LaunchTileForCout(float const* in, float const* w, float const* bias, OutT* out,
                       ConvDesc const& d, cudaStream_t stream)
{
    if (d.Cout <= 4 && KS == 3) LaunchTile<KS, 16, 64, 4, 1, CI_T>(in, w, bias, out, d, stream); // 4 px x 4 ch
    else if (d.Cout <= 4) LaunchTile<KS, 16, 16, 1, 1, CI_T>(in, w, bias, out, d, stream);
    else if (d.Cout <= 16) LaunchTile<KS, 16, 32, 8, 4, CI_T>(in, w, bias, out, d, stream);
    else if (d.Cout <= 32) LaunchTile<KS, 16, 16, 8, 8, CI_T>(in, w, bias, out, d, stream);
    else if (KS == 3) LaunchTile<KS, 16, 16, 8, 8, CI_T, 8>(in, w, bias, out, d, stream); // 8 px x 8 ch per thread
    else LaunchTile<KS, 8, 16, 8, 16, CI_T>(in, w, bias, out, d, stream);
}

__global__ void bias_act_kernel(float const* __restrict__ in, float const* __restrict__ bias,
                                float* __restrict__ out, size_t n, int C, Act act, float alpha)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride)
    {
        float v = in[i];
        if (bias) v += bias[i % C];
        out[i] = apply_act(v, act, alpha);
    }
}

// no __restrict__: `out` may be `a` or `b` (in-place inference, or the residual pass after a tensor-core layer)
__global__ void add_act_kernel(float const* a, float const* b, float* out, size_t n, Act act, float alpha)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride) out[i] = apply_act(a[i] + b[i], act, alpha);
}

__global__ void concat_kernel(float const* __restrict__ in, float* __restrict__ out, size_t pixels,
                              int C, int Ctotal, int c0)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    size_t const n = pixels * C;
    for (; i < n; i += stride)
    {
        size_t p = i / C;
        int c = static_cast<int>(i % C);
        out[p * Ctotal + c0 + c] = in[i];
    }
}

__global__ void half_to_float_kernel(__half const* __restrict__ in, float* __restrict__ out, size_t n)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride) out[i] = __half2float(in[i]);
}

__global__ void float_to_half_kernel(float const* __restrict__ in, __half* __restrict__ out, size_t n)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride) out[i] = __float2half_rn(in[i]);
}

// fp32 [pixels][C] -> bf16 [pixels][16], channels >= C zero: the tensor-core layer's K-major rows
// are at least 16 channels (32 bytes, one 32B-swizzle row).
__global__ void pad_to_bf16x16_kernel(float const* __restrict__ in, uint4* __restrict__ out, size_t pixels, int C)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < pixels; i += stride)
    {
        float v[16];
#pragma unroll
        for (int c = 0; c < 16; c++) v[c] = c < C ? in[i * C + c] : 0.f;
        uint32_t w[8];
#pragma unroll
        for (int c = 0; c < 8; c++)
        {
            __nv_bfloat162 b = __floats2bfloat162_rn(v[2 * c], v[2 * c + 1]);
            w[c] = *reinterpret_cast<uint32_t*>(&b);
        }
        out[2 * i + 0] = make_uint4(w[0], w[1], w[2], w[3]);
        out[2 * i + 1] = make_uint4(w[4], w[5], w[6], w[7]);
    }
}

// One thread per (pixel, group of 8 channels): 8 fp32 in, 16 bytes of hi and 16 bytes of lo out.
__global__ void split_bf16_hilo_kernel(float const* __restrict__ in, uint4* __restrict__ out, size_t pixels, int C,
                                       int Cp)
{
    int const groups = Cp >> 3;
    size_t const total = pixels * static_cast<size_t>(groups);
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    bool const vec = (C & 3) == 0;
    for (; i < total; i += stride)
    {
        size_t const pix = i / groups;
        int const c0 = static_cast<int>(i - pix * groups) << 3;
        float v[8];
        float const* src = in + pix * C + c0;
        if (vec && c0 + 8 <= C)
        {
            float4 a = *reinterpret_cast<float4 const*>(src), b = *reinterpret_cast<float4 const*>(src + 4);
            v[0] = a.x; v[1] = a.y; v[2] = a.z; v[3] = a.w; v[4] = b.x; v[5] = b.y; v[6] = b.z; v[7] = b.w;
        }
        else
        {
#pragma unroll
            for (int c = 0; c < 8; c++) v[c] = c0 + c < C ? src[c] : 0.f;
        }
        uint32_t hi[4], lo[4];
#pragma unroll
        for (int c = 0; c < 4; c++)
        {
            __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * c], v[2 * c + 1]);
            float2 const hf = __bfloat1622float2(h);
            __nv_bfloat162 l = __floats2bfloat162_rn(v[2 * c] - hf.x, v[2 * c + 1] - hf.y);
            hi[c] = *reinterpret_cast<uint32_t*>(&h);
            lo[c] = *reinterpret_cast<uint32_t*>(&l);
        }
        // row of 2*Cp bf16 = 2*groups units of 16 bytes: hi units first, lo units after
        size_t const row = pix * static_cast<size_t>(2 * groups);
        out[row + (c0 >> 3)] = make_uint4(hi[0], hi[1], hi[2], hi[3]);
        out[row + groups + (c0 >> 3)] = make_uint4(lo[0], lo[1], lo[2], lo[3]);
    }
}

// ---------------------------------------------------------------------------
// The ops around the convolutions of an encoder / decoder graph (TensorFlow op definitions, NHWC, N = 1).
// One thread per output element, channel index fastest: neighbouring threads touch neighbouring addresses.
// ---------------------------------------------------------------------------
// out = act(in * scale[c] + bias[c]); scale / bias may be nullptr (1 / 0).  `out` may alias `in`.
__global__ void scale_bias_act_kernel(float const* in, float const* __restrict__ scale, float const* __restrict__ bias,
                                      float* out, size_t n, int C, Act act, float alpha)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride)
    {
        int const c = static_cast<int>(i % C);
        float v = in[i];
        if (scale) v = v * scale[c];            // separate multiply and add, like Mul followed by BiasAdd
        if (bias) v = v + bias[c];
        out[i] = apply_act(v, act, alpha);
    }
}

// mode 0: a + b, 1: a - b, 2: a * b.  `out` may alias an operand.
__global__ void binary_act_kernel(float const* a, float const* b, float* out, size_t n, int mode, Act act, float alpha)
{
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride)
    {
        float const x = a[i], y = b[i];
        out[i] = apply_act(mode == 0 ? x + y : mode == 1 ? x - y : x * y, act, alpha);
    }
}

// MaxPool / AvgPool, TensorFlow SAME / VALID.  AvgPool divides by the number of IN-BOUNDS taps (padding excluded).
__global__ void pool2d_kernel(float const* __restrict__ in, float* __restrict__ out, PoolDesc d)
{
    size_t const n = static_cast<size_t>(d.OH) * d.OW * d.C;
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride)
    {
        int const c = static_cast<int>(i % d.C);
        size_t const p = i / d.C;
        int const ox = static_cast<int>(p % d.OW), oy = static_cast<int>(p / d.OW);
        int const y0 = oy * d.stride - d.pad_top, x0 = ox * d.stride - d.pad_left;
        float acc = d.is_max ? -__int_as_float(0x7f800000) : 0.0f;
        int taps = 0;
        for (int ky = 0; ky < d.K; ky++)
        {
            int const y = y0 + ky;
            if (y < 0 || y >= d.H) continue;
            for (int kx = 0; kx < d.K; kx++)
            {
                int const x = x0 + kx;
                if (x < 0 || x >= d.W) continue;
                float const v = in[(static_cast<size_t>(y) * d.W + x) * d.C + c];
                acc = d.is_max ? fmaxf(acc, v) : acc + v;
                taps++;
            }
        }
        out[i] = d.is_max ? acc : __fdiv_rn(acc, static_cast<float>(taps > 0 ? taps : 1));
    }
}

// ResizeNearestNeighbor / ResizeBilinear (tensorflow/core/kernels/image/resize_*.cc semantics):
//   scale = align_corners && out > 1 ? (in - 1) / (out - 1) : in / out
//   nearest : src = min(align_corners ? roundf(s(dst)) : floorf(s(dst)), in - 1)
//   bilinear: in = s(dst); lower = max(floor(in), 0); upper = min(ceil(in), in - 1); lerp = in - floor(in)
//   s(dst) = half_pixel_centers ? (dst + 0.5) * scale (nearest) / (dst + 0.5) * scale - 0.5 (bilinear) : dst * scale
__global__ void resize_kernel(float const* __restrict__ in, float* __restrict__ out, ResizeDesc d)
{
    size_t const n = static_cast<size_t>(d.OH) * d.OW * d.C;
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride)
    {
        int const c = static_cast<int>(i % d.C);
        size_t const p = i / d.C;
        int const ox = static_cast<int>(p % d.OW), oy = static_cast<int>(p / d.OW);
        if (!d.bilinear)
        {
            float const fy = d.half_pixel ? (static_cast<float>(oy) + 0.5f) * d.scale_y : static_cast<float>(oy) * d.scale_y;
            float const fx = d.half_pixel ? (static_cast<float>(ox) + 0.5f) * d.scale_x : static_cast<float>(ox) * d.scale_x;
            int const y = min(static_cast<int>(d.align_corners ? roundf(fy) : floorf(fy)), d.H - 1);
            int const x = min(static_cast<int>(d.align_corners ? roundf(fx) : floorf(fx)), d.W - 1);
            out[i] = in[(static_cast<size_t>(max(y, 0)) * d.W + max(x, 0)) * d.C + c];
        }
        else
        {
            float const fy = d.half_pixel ? (static_cast<float>(oy) + 0.5f) * d.scale_y - 0.5f : static_cast<float>(oy) * d.scale_y;
            float const fx = d.half_pixel ? (static_cast<float>(ox) + 0.5f) * d.scale_x - 0.5f : static_cast<float>(ox) * d.scale_x;
            float const fy0 = floorf(fy), fx0 = floorf(fx);
            int const y_lo = max(static_cast<int>(fy0), 0), y_hi = min(static_cast<int>(ceilf(fy)), d.H - 1);
            int const x_lo = max(static_cast<int>(fx0), 0), x_hi = min(static_cast<int>(ceilf(fx)), d.W - 1);
            float const ly = fy - fy0, lx = fx - fx0;
            auto at = [&](int y, int x) { return in[(static_cast<size_t>(y) * d.W + x) * d.C + c]; };
            float const tl = at(y_lo, x_lo), tr = at(y_lo, x_hi), bl = at(y_hi, x_lo), br = at(y_hi, x_hi);
            float const top = tl + (tr - tl) * lx, bottom = bl + (br - bl) * lx;
            out[i] = top + (bottom - top) * ly;
        }
    }
}

// DepthToSpace (NHWC, DCR): out[h*bs + i][w*bs + j][c] = in[h][w][(i*bs + j)*Cout + c]
__global__ void depth_to_space_kernel(float const* __restrict__ in, float* __restrict__ out, int H, int W, int C, int bs)
{
    int const Co = C / (bs * bs), OW = W * bs;
    size_t const n = static_cast<size_t>(H) * W * C;
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride)
    {
        int const c = static_cast<int>(i % Co);
        size_t const p = i / Co;
        int const ox = static_cast<int>(p % OW), oy = static_cast<int>(p / OW);
        int const h = oy / bs, bi = oy - h * bs, w = ox / bs, bj = ox - w * bs;
        out[i] = in[(static_cast<size_t>(h) * W + w) * C + (bi * bs + bj) * Co + c];
    }
}

// Pad (CONSTANT, zeros) / MirrorPad (REFLECT: edge not repeated, SYMMETRIC: edge repeated) on H and W.
__global__ void pad2d_kernel(float const* __restrict__ in, float* __restrict__ out, int H, int W, int C, int top,
                             int left, int OH, int OW, int mode)
{
    size_t const n = static_cast<size_t>(OH) * OW * C;
    size_t i = static_cast<size_t>(blockIdx.x) * blockDim.x + threadIdx.x;
    size_t const stride = static_cast<size_t>(gridDim.x) * blockDim.x;
    for (; i < n; i += stride)
    {
        int const c = static_cast<int>(i % C);
        size_t const p = i / C;
        int x = static_cast<int>(p % OW) - left, y = static_cast<int>(p / OW) - top;
        if (mode == 0)
        {
            out[i] = (x < 0 || x >= W || y < 0 || y >= H) ? 0.0f : in[(static_cast<size_t>(y) * W + x) * C + c];
            continue;
        }
        int const off = mode == 1 ? 0 : 1;         // REFLECT: -1 -> 1 ; SYMMETRIC: -1 -> 0
        if (x < 0) x = -x - off; else if (x >= W) x = 2 * W - x - 2 + off;
        if (y < 0) y = -y - off; else if (y >= H) y = 2 * H - y - 2 + off;
        out[i] = in[(static_cast<size_t>(y) * W + x) * C + c];
    }
}

inline unsigned EltGrid(size_t n)
{
    size_t blocks = (n + 255) / 256;
    size_t const cap = 148u * 16u;
    return static_cast<unsigned>(blocks < 1 ? 1 : (blocks > cap ? cap : blocks));
}

} // namespace

void Conv2dFp32(float const* in, float const* w, float const* bias, float* out, ConvDesc const& d,
                cudaStream_t stream)
{
    bool const tiled = d.stride == 1 && d.dilation == 1 && d.KH == d.KW
                    && (d.KH == 1 || d.KH == 3 || d.KH == 5);
    if (tiled)
    {
        if (d.KH == 1) LaunchTileForCout<1, 8>(in, w, bias, out, d, stream);
        else if (d.KH == 3) LaunchTileForCout<3, 8>(in, w, bias, out, d, stream);
        else LaunchTileForCout<5, 4>(in, w, bias, out, d, stream);
    }
    else
    {
        size_t const total = static_cast<size_t>(d.OH) * d.OW * d.Cout;
        unsigned const blocks = static_cast<unsigned>((total + 255) / 256);
        conv2d_naive_kernel<<<blocks, 256, 0, stream>>>(in, w, bias, out, d);
    }
    ML_CUDA_CHECK(cudaGetLastError());
}

void Conv2dFp32ToBf16(float const* in, float const* w, float const* bias, void* out_bf16,
                      ConvDesc const& d, cudaStream_t stream)
{
    // only the 3x3 stride-1 shape the tensor-core chain starts with
    if (!(d.stride == 1 && d.dilation == 1 && d.KH == 3 && d.KW == 3))
        throw std::runtime_error("Conv2dFp32ToBf16: only 3x3 stride 1");
    LaunchTileForCout<3, 8>(in, w, bias, static_cast<__nv_bfloat16*>(out_bf16), d, stream);
    ML_CUDA_CHECK(cudaGetLastError());
}

void PadToBf16x16(float const* in, void* out_bf16, size_t pixels, int C, cudaStream_t stream)
{
    if (pixels == 0) return;
    if (C < 1 || C > 16) throw std::runtime_error("PadToBf16x16: 1..16 channels");
    pad_to_bf16x16_kernel<<<EltGrid(pixels), 256, 0, stream>>>(in, static_cast<uint4*>(out_bf16), pixels, C);
    ML_CUDA_CHECK(cudaGetLastError());
}

void SplitToBf16HiLo(float const* in, void* out_bf16, size_t pixels, int C, int Cp, cudaStream_t stream)
{
    if (pixels == 0) return;
    if (C < 1 || Cp < C || (Cp & 7) != 0) throw std::runtime_error("SplitToBf16HiLo: bad channel counts");
    split_bf16_hilo_kernel<<<EltGrid(pixels * static_cast<size_t>(Cp >> 3)), 256, 0, stream>>>(
        in, static_cast<uint4*>(out_bf16), pixels, C, Cp);
    ML_CUDA_CHECK(cudaGetLastError());
}

void BiasAct(float const* in, float const* bias, float* out, size_t n, int C, Act act, float alpha,
             cudaStream_t stream)
{
    if (n == 0) return;
    bias_act_kernel<<<EltGrid(n), 256, 0, stream>>>(in, bias, out, n, C, act, alpha);
    ML_CUDA_CHECK(cudaGetLastError());
}

void AddAct(float const* a, float const* b, float* out, size_t n, Act act, float alpha,
            cudaStream_t stream)
{
    if (n == 0) return;
    add_act_kernel<<<EltGrid(n), 256, 0, stream>>>(a, b, out, n, act, alpha);
    ML_CUDA_CHECK(cudaGetLastError());
}

void ConcatChannels(float const* in, float* out, size_t pixels, int C, int Ctotal, int c0,
                    cudaStream_t stream)
{
    if (pixels == 0) return;
    concat_kernel<<<EltGrid(pixels * C), 256, 0, stream>>>(in, out, pixels, C, Ctotal, c0);
    ML_CUDA_CHECK(cudaGetLastError());
}

void ScaleBiasAct(float const* in, float const* scale, float const* bias, float* out, size_t n, int C, Act act,
                  float alpha, cudaStream_t stream)
{
    if (n == 0) return;
    scale_bias_act_kernel<<<EltGrid(n), 256, 0, stream>>>(in, scale, bias, out, n, C, act, alpha);
    ML_CUDA_CHECK(cudaGetLastError());
}

void BinaryAct(float const* a, float const* b, float* out, size_t n, int mode, Act act, float alpha, cudaStream_t stream)
{
    if (n == 0) return;
    binary_act_kernel<<<EltGrid(n), 256, 0, stream>>>(a, b, out, n, mode, act, alpha);
    ML_CUDA_CHECK(cudaGetLastError());
}

void Pool2d(float const* in, float* out, PoolDesc const& d, cudaStream_t stream)
{
    size_t const n = static_cast<size_t>(d.OH) * d.OW * d.C;
    if (n == 0) return;
    pool2d_kernel<<<EltGrid(n), 256, 0, stream>>>(in, out, d);
    ML_CUDA_CHECK(cudaGetLastError());
}

void Resize(float const* in, float* out, ResizeDesc const& d, cudaStream_t stream)
{
    size_t const n = static_cast<size_t>(d.OH) * d.OW * d.C;
    if (n == 0) return;
    resize_kernel<<<EltGrid(n), 256, 0, stream>>>(in, out, d);
    ML_CUDA_CHECK(cudaGetLastError());
}

void DepthToSpace(float const* in, float* out, int H, int W, int C, int block, cudaStream_t stream)
{
    size_t const n = static_cast<size_t>(H) * W * C;
    if (n == 0) return;
    depth_to_space_kernel<<<EltGrid(n), 256, 0, stream>>>(in, out, H, W, C, block);
    ML_CUDA_CHECK(cudaGetLastError());
}

void Pad2d(float const* in, float* out, int H, int W, int C, int top, int left, int OH, int OW, int mode,
           cudaStream_t stream)
{
    size_t const n = static_cast<size_t>(OH) * OW * C;
    if (n == 0) return;
    pad2d_kernel<<<EltGrid(n), 256, 0, stream>>>(in, out, H, W, C, top, left, OH, OW, mode);
    ML_CUDA_CHECK(cudaGetLastError());
}

void HalfToFloat(void const* in, float* out, size_t n, cudaStream_t stream)
{
    if (n == 0) return;
    half_to_float_kernel<<<EltGrid(n), 256, 0, stream>>>(static_cast<__half const*>(in), out, n);
    ML_CUDA_CHECK(cudaGetLastError());
}

void FloatToHalf(float const* in, void* out, size_t n, cudaStream_t stream)
{
    if (n == 0) return;
    float_to_half_kernel<<<EltGrid(n), 256, 0, stream>>>(in, static_cast<__half*>(out), n);
    ML_CUDA_CHECK(cudaGetLastError());
}

} // namespace nn
} // namespace ML
