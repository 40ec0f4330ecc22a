// conv_oracle.cpp -- CPU restatement of the inference arithmetic.  TEST
// INFRASTRUCTURE ONLY (tests/, __graft_entry__.smoke(), bench.py CPU-baseline
// legs); nothing in simpleraytracer_b200/ links or calls it.
//
// PARITY UNPINNED: in the reference all arithmetic is one call into TensorFlow,
// tensorflow::Session::Run at model_runner/model.cpp:339.  TensorFlow is an
// un-vendored, unpinned third-party dependency (README suggests "e.g. v1.14",
// model_runner/README.md:124), no model file and no golden vectors are shipped
// (SURVEY.md section 8c).  What is restated here is the published definition of the
// TensorFlow ops a frozen image-to-image graph uses:
//   Conv2D (NHWC input, HWIO filter, SAME/VALID padding, stride, dilation),
//   BiasAdd, Relu / LeakyRelu / Relu6 / Sigmoid / Tanh, Add, ConcatV2(axis=3),
//   MaxPool / AvgPool, ResizeNearestNeighbor / ResizeBilinear (align_corners, half_pixel_centers),
//   DepthToSpace, Pad / MirrorPad, FusedBatchNorm (inference), Mul / Sub / RealDiv -- the ops around the
//   convolutions of an encoder / decoder denoiser (tensorflow/core/kernels/{pooling_ops_common,
//   image/resize_*_op, depthtospace_op, pad_op, mirror_pad_op}: published semantics, restated scalar).
// Fixed accumulation order (kh, kw, cin), fp32 with fmaf.  Independent
// cross-check in tests: torch.nn.functional.conv2d on the CPU.
#include <algorithm>
#include <cmath>
#include <cstdint>
#include <thread>
#include <vector>

namespace {

inline float Activate(float v, int act, float alpha)
{
    switch (act)
    {
        case 1: return v > 0.f ? v : 0.f;                       // Relu
        case 2: return v > 0.f ? v : v * alpha;                 // LeakyRelu
        case 3: return v < 0.f ? 0.f : (v > 6.f ? 6.f : v);     // Relu6
        case 4: return 1.0f / (1.0f + std::exp(-v));            // Sigmoid
        case 5: return std::tanh(v);                            // Tanh
        default: return v;
    }
}

template <class Fn>
void ParallelRows(int rows, int threads, Fn fn)
{
    int n = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    n = std::max(1, std::min(n, rows));
    std::vector<std::thread> pool;
    for (int t = 0; t < n; t++)
        pool.emplace_back([=]() {
            for (int r = t; r < rows; r += n) fn(r);
        });
    for (auto& th : pool) th.join();
}

} // namespace

extern "C" {

// TensorFlow output size / low padding for one spatial dimension.
void orc_conv_geometry(int in, int k, int stride, int dilation, int same, int* out, int* pad_lo)
{
    int eff = (k - 1) * dilation + 1;
    if (same)
    {
        *out = (in + stride - 1) / stride;
        int total = std::max((*out - 1) * stride + eff - in, 0);
        *pad_lo = total / 2;
    }
    else
    {
        *out = in >= eff ? (in - eff) / stride + 1 : 0;
        *pad_lo = 0;
    }
}

// in: H*W*Cin; w: KH*KW*Cin*Cout (HWIO); bias: Cout or NULL; out: OH*OW*Cout.
// act: 0 none, 1 relu, 2 leaky, 3 relu6, 4 sigmoid, 5 tanh.  Returns threads used.
int orc_conv2d(float const* in, int H, int W, int Cin, float const* w, int KH, int KW, int Cout,
               float const* bias, int stride, int dilation, int same, int act, float alpha,
               float* out, int threads)
{
    int OH, OW, pt, pl;
    orc_conv_geometry(H, KH, stride, dilation, same, &OH, &pt);
    orc_conv_geometry(W, KW, stride, dilation, same, &OW, &pl);
    ParallelRows(OH, threads, [=](int oy) {
        std::vector<float> acc(static_cast<size_t>(Cout));
        for (int ox = 0; ox < OW; ox++)
        {
            std::fill(acc.begin(), acc.end(), 0.f);
            for (int kh = 0; kh < KH; kh++)
            {
                int y = oy * stride + kh * dilation - pt;
                if (y < 0 || y >= H) continue;
                for (int kw = 0; kw < KW; kw++)
                {
                    int x = ox * stride + kw * dilation - pl;
                    if (x < 0 || x >= W) continue;
                    float const* ip = in + (static_cast<size_t>(y) * W + x) * Cin;
                    float const* wp = w + (static_cast<size_t>(kh) * KW + kw) * Cin * Cout;
                    for (int ci = 0; ci < Cin; ci++)
                    {
                        float v = ip[ci];
                        float const* wr = wp + static_cast<size_t>(ci) * Cout;
                        for (int co = 0; co < Cout; co++) acc[co] = std::fmaf(v, wr[co], acc[co]);
                    }
                }
            }
            float* op = out + (static_cast<size_t>(oy) * OW + ox) * Cout;
            for (int co = 0; co < Cout; co++)
                op[co] = Activate(acc[co] + (bias ? bias[co] : 0.f), act, alpha);
        }
    });
    int n = threads > 0 ? threads : static_cast<int>(std::thread::hardware_concurrency());
    return std::max(1, std::min(n, OH));
}

// out[i] = act(a[i] + (b ? b[i] : 0) + (bias ? bias[i % C] : 0))
void orc_add_bias_act(float const* a, float const* b, float const* bias, uint64_t n, int C, int act,
                      float alpha, float* out)
{
    for (uint64_t i = 0; i < n; i++)
    {
        float v = a[i];
        if (b) v = v + b[i];
        if (bias) v = v + bias[i % static_cast<uint64_t>(C)];
        out[i] = Activate(v, act, alpha);
    }
}

// MaxPool (is_max) / AvgPool over K x K windows, stride, TensorFlow SAME / VALID; AvgPool divides by the number of
// in-bounds taps.  out: OH x OW x C with OH/OW from orc_conv_geometry(in, K, stride, 1, same).
void orc_pool2d(float const* in, int H, int W, int C, int K, int stride, int same, int is_max, float* out)
{
    int OH, OW, pt, pl;
    orc_conv_geometry(H, K, stride, 1, same, &OH, &pt);
    orc_conv_geometry(W, K, stride, 1, same, &OW, &pl);
    for (int oy = 0; oy < OH; oy++)
        for (int ox = 0; ox < OW; ox++)
            for (int c = 0; c < C; c++)
            {
                float acc = is_max ? -INFINITY : 0.f;
                int taps = 0;
                for (int ky = 0; ky < K; ky++)
                    for (int kx = 0; kx < K; kx++)
                    {
                        int const y = oy * stride - pt + ky, x = ox * stride - pl + kx;
                        if (y < 0 || y >= H || x < 0 || x >= W) continue;
                        float const v = in[(static_cast<size_t>(y) * W + x) * C + c];
                        acc = is_max ? std::max(acc, v) : acc + v;
                        taps++;
                    }
                out[(static_cast<size_t>(oy) * OW + ox) * C + c] = is_max ? acc : acc / static_cast<float>(taps ? taps : 1);
            }
}

// ResizeNearestNeighbor (bilinear = 0) / ResizeBilinear (1): tensorflow/core/kernels/image/resize_*_op.cc and
// image_resizer_state.h (CalculateResizeScale, LegacyScaler / HalfPixelScaler, compute_interpolation_weights).
void orc_resize(float const* in, int H, int W, int C, int OH, int OW, int bilinear, int align_corners, int half_pixel,
                float* out)
{
    auto scale = [&](int i, int o) {
        return (align_corners && o > 1) ? static_cast<float>(i - 1) / static_cast<float>(o - 1)
                                        : static_cast<float>(i) / static_cast<float>(o);
    };
    float const sy = scale(H, OH), sx = scale(W, OW);
    for (int oy = 0; oy < OH; oy++)
        for (int ox = 0; ox < OW; ox++)
            for (int c = 0; c < C; c++)
            {
                float r;
                if (!bilinear)
                {
                    float const fy = half_pixel ? (static_cast<float>(oy) + 0.5f) * sy : static_cast<float>(oy) * sy;
                    float const fx = half_pixel ? (static_cast<float>(ox) + 0.5f) * sx : static_cast<float>(ox) * sx;
                    int y = std::min(static_cast<int>(align_corners ? std::round(fy) : std::floor(fy)), H - 1);
                    int x = std::min(static_cast<int>(align_corners ? std::round(fx) : std::floor(fx)), W - 1);
                    y = std::max(y, 0); x = std::max(x, 0);
                    r = in[(static_cast<size_t>(y) * W + x) * C + c];
                }
                else
                {
                    float const fy = half_pixel ? (static_cast<float>(oy) + 0.5f) * sy - 0.5f : static_cast<float>(oy) * sy;
                    float const fx = half_pixel ? (static_cast<float>(ox) + 0.5f) * sx - 0.5f : static_cast<float>(ox) * sx;
                    float const fy0 = std::floor(fy), fx0 = std::floor(fx);
                    int const y0 = std::max(static_cast<int>(fy0), 0), y1 = std::min(static_cast<int>(std::ceil(fy)), H - 1);
                    int const x0 = std::max(static_cast<int>(fx0), 0), x1 = std::min(static_cast<int>(std::ceil(fx)), W - 1);
                    float const ly = fy - fy0, lx = fx - fx0;
                    auto at = [&](int y, int x) { return in[(static_cast<size_t>(y) * W + x) * C + c]; };
                    float const top = at(y0, x0) + (at(y0, x1) - at(y0, x0)) * lx;
                    float const bottom = at(y1, x0) + (at(y1, x1) - at(y1, x0)) * lx;
                    r = top + (bottom - top) * ly;
                }
                out[(static_cast<size_t>(oy) * OW + ox) * C + c] = r;
            }
}

// DepthToSpace, NHWC: out[h*bs + i][w*bs + j][c] = in[h][w][(i*bs + j)*Co + c], Co = C / bs^2.
void orc_depth_to_space(float const* in, int H, int W, int C, int bs, float* out)
{
    int const Co = C / (bs * bs), OW = W * bs;
    for (int h = 0; h < H; h++)
        for (int w = 0; w < W; w++)
            for (int i = 0; i < bs; i++)
                for (int j = 0; j < bs; j++)
                    for (int c = 0; c < Co; c++)
                        out[((static_cast<size_t>(h) * bs + i) * OW + (w * bs + j)) * Co + c] =
                            in[(static_cast<size_t>(h) * W + w) * C + (i * bs + j) * Co + c];
}

// Pad (mode 0: zeros) / MirrorPad (1: REFLECT, 2: SYMMETRIC) on the H and W axes.
void orc_pad2d(float const* in, int H, int W, int C, int top, int bottom, int left, int right, int mode, float* out)
{
    int const OH = H + top + bottom, OW = W + left + right;
    auto mirror = [&](int v, int n) {
        int const off = mode == 1 ? 0 : 1;
        if (v < 0) return -v - off;
        if (v >= n) return 2 * n - v - 2 + off;
        return v;
    };
    for (int oy = 0; oy < OH; oy++)
        for (int ox = 0; ox < OW; ox++)
        {
            int y = oy - top, x = ox - left;
            bool const outside = y < 0 || y >= H || x < 0 || x >= W;
            if (mode != 0) { y = mirror(y, H); x = mirror(x, W); }
            for (int c = 0; c < C; c++)
                out[(static_cast<size_t>(oy) * OW + ox) * C + c] =
                    (mode == 0 && outside) ? 0.f : in[(static_cast<size_t>(y) * W + x) * C + c];
        }
}

} // extern "C"
