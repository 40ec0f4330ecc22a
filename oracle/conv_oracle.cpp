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
//   BiasAdd, Relu / LeakyRelu / Relu6 / Sigmoid / Tanh, Add, ConcatV2(axis=3).
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

} // extern "C"
