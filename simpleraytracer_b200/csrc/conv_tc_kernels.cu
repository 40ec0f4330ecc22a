// conv_tc_kernels.cu -- 3x3 / stride 1 / SAME convolution in bf16 on the 5th-gen
// tensor cores (tcgen05.mma, fp32 accumulators in TMEM), for the wide layers of
// the frozen graph that stood behind tensorflow::Session::Run (reference
// model.cpp:339).  Hand-written sm_100a: TMA (cp.async.bulk.tensor) stages the
// operands, one elected thread issues the MMAs, four warps drain TMEM.
//
// Implicit GEMM, no im2col buffer.  For an output tile of 128 consecutive
// pixels of one image row (M = 128), N = Cout (padded to 16) and K = 9 * Cin:
//     D[pixel, co] = sum over the 9 taps (kh, kw) of  A_tap[pixel, ci] * B_tap[co, ci]
// where A_tap is the NHWC activation row y+kh-1 shifted by kw-1 pixels -- a box
// of the activation tensor that TMA fetches directly, zero-filling whatever
// lies outside the image (that IS the SAME padding) -- and B_tap the filter
// slice repacked once on the host as [tap][co][ci] (K-major).  Both operands are
// K-major in shared memory under the 128/64/32-byte swizzle that matches the
// row length Cin * 2 bytes, so the TMA box lands in exactly the canonical
// layout the UMMA shared-memory descriptor describes.
//
// Persistent kernel, one CTA per SM, 6 warps:
//   warp 0 (one lane)   TMA producer: the filter (once, stays resident in smem),
//                       then a ring of A stages, one per (tile, tap, Cin chunk)
//   warp 1 (one lane)   MMA issuer: Cin/16 tcgen05.mma per stage, tcgen05.commit
//                       frees the stage / publishes the accumulator
//   warps 2-5           epilogue: tcgen05.ld 32 lanes x 16 columns at a time,
//                       + bias, activation, pack to bf16 (or fp32 for the last
//                       layer), 16-byte stores; the accumulator is double
//                       buffered in TMEM so tile i's epilogue overlaps tile
//                       i+1's MMAs.
#include "nn_kernels.h"

#include "common.h"

#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_bf16.h>

#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace ML {
namespace nn {
namespace {

constexpr int kTileM = 128;          // output pixels per tile = TMEM lanes
constexpr int kTcThreads = 192;      // 6 warps
constexpr int kMaxStages = 8;

struct TcParams
{
    int H, W;            // image
    int Cin, Cout;       // real channel counts
    int N;               // Cout padded to a multiple of 16 (MMA N)
    int KC;              // Cin chunk per stage (<= 64): one swizzle row
    int n_chunks;        // Cin / KC
    int stages;          // A ring depth
    int tiles_per_row;   // ceil(W / 128)
    int n_tiles;
    int layout_type;     // UMMA LayoutType: 2 = SW128, 4 = SW64, 6 = SW32
    int sbo;             // bytes between 8-row groups = 8 * KC * 2
    int tmem_cols;       // allocated TMEM columns (power of two >= 32, holds 2 accumulators)
    int out_bf16;        // 1: bf16 NHWC with N == Cout channels; 0: fp32 NHWC with Cout channels
    Act act;
    float alpha;
    int halo;            // 1: a stage is ONE box of 130 pixels (row y+kh-1, x0-1..x0+128) and the three
                         //    kw taps are row-shifted views of it (KC == 64 only); 0: three 128-pixel boxes
    int stage_bytes;     // bytes of one A stage (one kh row of one Cin chunk), multiple of 1024
    int tap_stride;      // bytes from tap kw to tap kw+1 inside a stage: 128 (halo) or 128*KC*2
    int stage_tx;        // bytes TMA delivers per stage
    int base_offset_mode; // halo mode: fill the descriptor's base-offset field (MLB200_TC_BASEOFF, default 0: not needed)
    int debug;           // MLB200_TC_DEBUG bit 0: no TMA loads of A, bit 1: no MMAs (timing experiments only)
};

// ---------------------------------------------------------------- PTX helpers
__device__ __forceinline__ uint32_t smem_addr(void const* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void bar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_addr(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void bar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_addr(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void bar_arrive(uint64_t* bar)
{
    asm volatile("mbarrier.arrive.shared::cta.b64 _, [%0];" ::"r"(smem_addr(bar)) : "memory");
}
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    do
    {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void tma_load_3d(void* dst, CUtensorMap const* map, int c0, int c1, int c2,
                                            uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%2, %3, %4}], [%5];"
        ::"r"(smem_addr(dst)), "l"(map), "r"(c0), "r"(c1), "r"(c2), "r"(smem_addr(bar))
        : "memory");
}
__device__ __forceinline__ void tma_load_2d(void* dst, CUtensorMap const* map, int c0, int c1, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.tensor.2d.shared::cluster.global.tile.mbarrier::complete_tx::bytes"
        " [%0], [%1, {%2, %3}], [%4];"
        ::"r"(smem_addr(dst)), "l"(map), "r"(c0), "r"(c1), "r"(smem_addr(bar))
        : "memory");
}
__device__ __forceinline__ void tc_fence_before() { asm volatile("tcgen05.fence::before_thread_sync;" ::: "memory"); }
__device__ __forceinline__ void tc_fence_after() { asm volatile("tcgen05.fence::after_thread_sync;" ::: "memory"); }

// tcgen05.commit: the mbarrier receives one arrival when every MMA issued so far
// by this thread has finished reading shared memory / writing TMEM.
__device__ __forceinline__ void tc_commit(uint64_t* bar)
{
    asm volatile("tcgen05.commit.cta_group::1.mbarrier::arrive::one.shared::cluster.b64 [%0];"
                 ::"r"(smem_addr(bar)) : "memory");
}
// D[tmem] (+)= A[smem] * B[smem], bf16 x bf16 -> fp32, issued by ONE thread for the CTA.
__device__ __forceinline__ void tc_mma_bf16(uint32_t d_tmem, uint64_t a_desc, uint64_t b_desc, uint32_t idesc,
                                            uint32_t accumulate)
{
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "setp.ne.b32 p, %4, 0;\n"
        "tcgen05.mma.cta_group::1.kind::f16 [%0], %1, %2, %3, p;\n"
        "}\n"
        ::"r"(d_tmem), "l"(a_desc), "l"(b_desc), "r"(idesc), "r"(accumulate)
        : "memory");
}
// 32 lanes x 16 consecutive fp32 columns of TMEM -> 16 registers per thread.
__device__ __forceinline__ void tc_ld16(uint32_t taddr, uint32_t (&r)[16])
{
    asm volatile(
        "tcgen05.ld.sync.aligned.32x32b.x16.b32 "
        "{%0, %1, %2, %3, %4, %5, %6, %7, %8, %9, %10, %11, %12, %13, %14, %15}, [%16];"
        : "=r"(r[0]), "=r"(r[1]), "=r"(r[2]), "=r"(r[3]), "=r"(r[4]), "=r"(r[5]), "=r"(r[6]), "=r"(r[7]),
          "=r"(r[8]), "=r"(r[9]), "=r"(r[10]), "=r"(r[11]), "=r"(r[12]), "=r"(r[13]), "=r"(r[14]), "=r"(r[15])
        : "r"(taddr)
        : "memory");
}
__device__ __forceinline__ void tc_ld_wait() { asm volatile("tcgen05.wait::ld.sync.aligned;" ::: "memory"); }

// K-major operand under a 128/64/32-byte swizzle: rows of KC*2 bytes, 8-row groups
// `sbo` bytes apart (cute::UMMA::SmemDescriptor bit layout: start>>4 [0,14), LBO>>4
// [16,30), SBO>>4 [32,46), version=1 [46,48), layout type [61,64)).
__device__ __forceinline__ uint64_t umma_desc(uint32_t saddr, uint32_t sbo, uint32_t layout_type)
{
    uint64_t d = 0;
    d |= static_cast<uint64_t>((saddr >> 4) & 0x3FFFu);
    d |= static_cast<uint64_t>(1u) << 16;                       // LBO: unused for swizzled K-major
    d |= static_cast<uint64_t>((sbo >> 4) & 0x3FFFu) << 32;
    d |= static_cast<uint64_t>(1u) << 46;                       // descriptor version (Blackwell)
    d |= static_cast<uint64_t>(layout_type & 7u) << 61;
    return d;
}

__device__ __forceinline__ float tc_act(float v, Act act, float alpha)
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
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi)
{
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// ---------------------------------------------------------------- the kernel
__global__ void __launch_bounds__(kTcThreads, 1)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap w_map,
                  float const* __restrict__ bias, void* __restrict__ out, TcParams const P)
{
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment: the 128-byte swizzle pattern repeats every 1024 bytes
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~uintptr_t(1023u));
    uint32_t const b_tile_bytes = static_cast<uint32_t>(P.N) * P.KC * 2u;
    uint32_t const a_stage_bytes = static_cast<uint32_t>(P.stage_bytes);
    uint32_t const n_b_tiles = 9u * P.n_chunks;
    uint8_t* smem_b = smem;
    uint8_t* smem_a = smem + n_b_tiles * b_tile_bytes;
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_a + static_cast<uint32_t>(P.stages) * a_stage_bytes);
    uint64_t* full = bars;                       // [stages]   TMA -> MMA
    uint64_t* empty = bars + kMaxStages;         // [stages]   MMA -> TMA
    uint64_t* w_full = bars + 2 * kMaxStages;    // filter resident
    uint64_t* acc_full = w_full + 1;             // [2]        MMA -> epilogue
    uint64_t* acc_empty = acc_full + 2;          // [2]        epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + 2);

    uint32_t const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // bias (zero beyond Cout) staged once; the epilogue reads it as broadcast LDS.128
    __shared__ __align__(16) float s_bias[256];
    for (int c = threadIdx.x; c < P.N; c += kTcThreads) s_bias[c] = (bias != nullptr && c < P.Cout) ? bias[c] : 0.f;
    bool const simple_act = P.act == Act::None || P.act == Act::Relu || P.act == Act::LeakyRelu;
    float const neg_slope = P.act == Act::None ? 1.f : (P.act == Act::LeakyRelu ? P.alpha : 0.f);

    if (warp == 0 && lane == 0)
    {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&in_map) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&w_map) : "memory");
        for (int s = 0; s < P.stages; s++) { bar_init(&full[s], 1); bar_init(&empty[s], 1); }
        bar_init(w_full, 1);
        for (int a = 0; a < 2; a++) { bar_init(&acc_full[a], 1); bar_init(&acc_empty[a], 4); }
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    if (warp == 1)
    {
        asm volatile("tcgen05.alloc.cta_group::1.sync.aligned.shared::cta.b32 [%0], %1;"
                     ::"r"(smem_addr(tmem_slot)), "r"(static_cast<uint32_t>(P.tmem_cols)) : "memory");
        asm volatile("tcgen05.relinquish_alloc_permit.cta_group::1.sync.aligned;" ::: "memory");
    }
    tc_fence_before();
    __syncthreads();
    tc_fence_after();
    uint32_t const tmem_base = *tmem_slot;

    if (warp == 0)
    {
        if (lane == 0)
        {
            // ---- TMA producer
            bar_expect_tx(w_full, n_b_tiles * b_tile_bytes);
            for (uint32_t t = 0; t < 9u; t++)
                for (int ch = 0; ch < P.n_chunks; ch++)
                    tma_load_2d(smem_b + (t * P.n_chunks + ch) * b_tile_bytes, &w_map, ch * P.KC,
                                static_cast<int>(t) * P.N, w_full);
            uint32_t it = 0;
            for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x)
            {
                int const y = tile / P.tiles_per_row;
                int const x0 = (tile - y * P.tiles_per_row) * kTileM;
                for (int kh = 0; kh < 3; kh++)
                {
                    for (int ch = 0; ch < P.n_chunks; ch++, it++)
                    {
                        uint32_t const s = it % P.stages, ph = (it / P.stages) & 1u;
                        bar_wait(&empty[s], ph ^ 1u);
                        if (P.debug & 1) { bar_arrive(&full[s]); continue; }
                        bar_expect_tx(&full[s], static_cast<uint32_t>(P.stage_tx));
                        // rows/pixels outside the image are zero-filled by TMA = SAME padding
                        uint8_t* dst = smem_a + s * a_stage_bytes;
                        if (P.halo)
                        {
                            tma_load_3d(dst, &in_map, ch * P.KC, x0 - 1, y + kh - 1, &full[s]);
                        }
                        else
                        {
                            for (int kw = 0; kw < 3; kw++)
                                tma_load_3d(dst + kw * P.tap_stride, &in_map, ch * P.KC, x0 + kw - 1, y + kh - 1,
                                            &full[s]);
                        }
                    }
                }
            }
        }
    }
    else if (warp == 1)
    {
        if (lane == 0)
        {
            // ---- MMA issuer.  Instruction descriptor (cute::UMMA::InstrDescriptor): D = fp32 [4,6),
            // A = B = bf16 [7,10) [10,13), both K-major, N>>3 at [17,23), M>>4 at [24,29).
            uint32_t const idesc = (1u << 4) | (1u << 7) | (1u << 10)
                                 | (static_cast<uint32_t>(P.N >> 3) << 17)
                                 | (static_cast<uint32_t>(kTileM >> 4) << 24);
            bar_wait(w_full, 0);
            tc_fence_after();
            uint32_t it = 0, tcount = 0;
            for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, tcount++)
            {
                uint32_t const acc = tcount & 1u, aph = (tcount >> 1) & 1u;
                bar_wait(&acc_empty[acc], aph ^ 1u);
                tc_fence_after();
                uint32_t const d_tmem = tmem_base + acc * static_cast<uint32_t>(P.N);
                uint32_t first = 0;
                for (uint32_t kh = 0; kh < 3u; kh++)
                {
                    for (int ch = 0; ch < P.n_chunks; ch++, it++)
                    {
                        uint32_t const s = it % P.stages, ph = (it / P.stages) & 1u;
                        bar_wait(&full[s], ph);
                        tc_fence_after();
                        uint32_t const a_base = smem_addr(smem_a + s * a_stage_bytes);
                        for (uint32_t kw = 0; kw < 3u; kw++)
                        {
                            // halo mode: tap kw = the same box read from pixel row kw on.  The start is then
                            // kw * 128 bytes into the 1024-byte swizzle pattern.  Measured on B200: the MMA
                            // applies the 128B swizzle XOR to the ABSOLUTE shared-memory address bits (as TMA
                            // does), so a row-shifted start needs NO base offset (bits [49,52) stay 0; setting
                            // them to (start >> 7) & 7 gives wrong results).  MLB200_TC_BASEOFF=1 re-enables
                            // the field for that experiment.
                            uint32_t const a_addr = a_base + kw * static_cast<uint32_t>(P.tap_stride);
                            uint64_t a_desc = umma_desc(a_addr, P.sbo, P.layout_type);
                            if (P.halo && P.base_offset_mode) a_desc |= static_cast<uint64_t>((a_addr >> 7) & 7u) << 49;
                            uint64_t const b_desc =
                                umma_desc(smem_addr(smem_b + ((kh * 3u + kw) * P.n_chunks + ch) * b_tile_bytes), P.sbo,
                                          P.layout_type);
                            for (int k = 0; k < P.KC / 16 && !(P.debug & 2); k++)
                            {
                                // +32 bytes along K inside the swizzle row = +2 in the (addr >> 4) field
                                tc_mma_bf16(d_tmem, a_desc + 2u * k, b_desc + 2u * k, idesc, first);
                                first = 1;
                            }
                        }
                        tc_commit(&empty[s]);       // ONE commit per stage (12 MMAs at Cin = 64)
                    }
                }
                tc_commit(&acc_full[acc]);          // accumulator complete
            }
        }
    }
    else
    {
        // ---- epilogue warps 2..5: warp w may touch TMEM lanes 32*(w%4) .. +31
        uint32_t const q = warp & 3u;
        uint32_t const row = q * 32u + lane;
        uint32_t tcount = 0;
        for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x, tcount++)
        {
            uint32_t const acc = tcount & 1u, aph = (tcount >> 1) & 1u;
            int const y = tile / P.tiles_per_row;
            int const x = (tile - y * P.tiles_per_row) * kTileM + static_cast<int>(row);
            bar_wait(&acc_full[acc], aph);
            tc_fence_after();
            uint32_t const taddr = tmem_base + acc * static_cast<uint32_t>(P.N) + ((q * 32u) << 16);
            size_t const pixel = static_cast<size_t>(y) * P.W + x;
            for (int c0 = 0; c0 < P.N; c0 += 32)
            {
                // two 16-column loads in flight, one wait (N is a multiple of 16: the second
                // half is warp-uniformly skipped on the last odd chunk)
                uint32_t r[32];
                bool const two = c0 + 16 < P.N;
                tc_ld16(taddr + static_cast<uint32_t>(c0), *reinterpret_cast<uint32_t(*)[16]>(&r[0]));
                if (two) tc_ld16(taddr + static_cast<uint32_t>(c0 + 16), *reinterpret_cast<uint32_t(*)[16]>(&r[16]));
                tc_ld_wait();
#pragma unroll
                for (int h = 0; h < 2; h++)
                {
                    if (h == 1 && !two) break;
                    int const cb = c0 + 16 * h;
                    float v[16];
#pragma unroll
                    for (int j4 = 0; j4 < 4; j4++)
                    {
                        float4 const b = *reinterpret_cast<float4 const*>(&s_bias[cb + 4 * j4]); // broadcast LDS.128
                        v[4 * j4 + 0] = __uint_as_float(r[16 * h + 4 * j4 + 0]) + b.x;
                        v[4 * j4 + 1] = __uint_as_float(r[16 * h + 4 * j4 + 1]) + b.y;
                        v[4 * j4 + 2] = __uint_as_float(r[16 * h + 4 * j4 + 2]) + b.z;
                        v[4 * j4 + 3] = __uint_as_float(r[16 * h + 4 * j4 + 3]) + b.w;
                    }
                    if (simple_act)
                    {
                        // None / Relu / LeakyRelu without branches: max(v,0) + slope * min(v,0)
#pragma unroll
                        for (int j = 0; j < 16; j++) v[j] = fmaf(neg_slope, fminf(v[j], 0.f), fmaxf(v[j], 0.f));
                    }
                    else
                    {
#pragma unroll
                        for (int j = 0; j < 16; j++) v[j] = tc_act(v[j], P.act, P.alpha);
                    }
                    if (x < P.W)
                    {
                        if (P.out_bf16)
                        {
                            uint4* o = reinterpret_cast<uint4*>(static_cast<__nv_bfloat16*>(out) + pixel * P.N + cb);
                            o[0] = make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                                              pack_bf16(v[6], v[7]));
                            o[1] = make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]),
                                              pack_bf16(v[12], v[13]), pack_bf16(v[14], v[15]));
                        }
                        else
                        {
                            float* o = static_cast<float*>(out) + pixel * P.Cout;
#pragma unroll
                            for (int j = 0; j < 16; j++)
                                if (cb + j < P.Cout) o[cb + j] = v[j];
                        }
                    }
                }
            }
            tc_fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(&acc_empty[acc]);
        }
    }

    tc_fence_before();
    __syncthreads();
    if (warp == 1)
    {
        tc_fence_after();
        asm volatile("tcgen05.dealloc.cta_group::1.sync.aligned.b32 %0, %1;"
                     ::"r"(tmem_base), "r"(static_cast<uint32_t>(P.tmem_cols)) : "memory");
    }
}

// ---------------------------------------------------------------- host side
PFN_cuTensorMapEncodeTiled_v12000 EncodeTiled()
{
    static PFN_cuTensorMapEncodeTiled_v12000 fn = [] {
        void* p = nullptr;
        cudaDriverEntryPointQueryResult q{};
        // The driver API entry point is fetched through the runtime: no link against libcuda.
        cudaError_t e = cudaGetDriverEntryPoint("cuTensorMapEncodeTiled", &p, cudaEnableDefault, &q);
        if (e != cudaSuccess || q != cudaDriverEntryPointSuccess || p == nullptr)
            throw std::runtime_error("cuTensorMapEncodeTiled is not available from this driver");
        return reinterpret_cast<PFN_cuTensorMapEncodeTiled_v12000>(p);
    }();
    return fn;
}

CUtensorMapSwizzle SwizzleFor(int kc)
{
    return kc == 64 ? CU_TENSOR_MAP_SWIZZLE_128B : kc == 32 ? CU_TENSOR_MAP_SWIZZLE_64B : CU_TENSOR_MAP_SWIZZLE_32B;
}

void Encode(CUtensorMap* map, void const* base, int rank, cuuint64_t const* dims, cuuint64_t const* strides,
            cuuint32_t const* box, CUtensorMapSwizzle swz)
{
    cuuint32_t estr[3] = {1, 1, 1};
    CUresult r = EncodeTiled()(map, CU_TENSOR_MAP_DATA_TYPE_BFLOAT16, static_cast<cuuint32_t>(rank),
                               const_cast<void*>(base), dims, strides, box, estr, CU_TENSOR_MAP_INTERLEAVE_NONE, swz,
                               CU_TENSOR_MAP_L2_PROMOTION_L2_256B, CU_TENSOR_MAP_FLOAT_OOB_FILL_NONE);
    if (r != CUDA_SUCCESS)
        throw std::runtime_error("cuTensorMapEncodeTiled failed with CUresult " + std::to_string(static_cast<int>(r)));
}

bool HaloMode(int KC)
{
    static int const env = std::getenv("MLB200_TC_HALO") ? std::atoi(std::getenv("MLB200_TC_HALO")) : 1;
    return KC == 64 && env != 0;
}

size_t StageBytes(int KC)
{
    // halo: 130 pixels x 128 B = 16640 B, rounded up to the 1024-byte swizzle pattern; else 3 boxes
    return HaloMode(KC) ? 17u * 1024u : static_cast<size_t>(3) * kTileM * KC * 2;
}

size_t SmemBytes(int N, int Cin, int KC, int stages)
{
    size_t b = static_cast<size_t>(9) * N * Cin * 2;
    size_t a = static_cast<size_t>(stages) * StageBytes(KC);
    return 1024 /*alignment slack*/ + b + a + (2 * kMaxStages + 5) * sizeof(uint64_t) + 16;
}

} // namespace

struct ConvTc
{
    CUtensorMap in_map;
    CUtensorMap w_map;
    TcParams p;
    size_t smem;
    int grid;
};

int ConvTcPaddedCout(int Cout) { return (Cout + 15) / 16 * 16; }

bool ConvTcSupported(int KH, int KW, int stride, int dilation, bool same_padding, int Cin, int Cout)
{
    if (KH != 3 || KW != 3 || stride != 1 || dilation != 1 || !same_padding) return false;
    if (!(Cin == 16 || Cin == 32 || (Cin > 0 && Cin % 64 == 0))) return false;
    int const N = ConvTcPaddedCout(Cout);
    if (Cout < 1 || N > 256) return false;
    int const KC = Cin < 64 ? Cin : 64;
    return SmemBytes(N, Cin, KC, 2) <= 226u * 1024u;
}

std::vector<uint16_t> ConvTcPackWeights(float const* hwio, int Cin, int Cout)
{
    int const N = ConvTcPaddedCout(Cout);
    std::vector<uint16_t> packed(static_cast<size_t>(9) * N * Cin, 0);
    for (int t = 0; t < 9; t++)
        for (int co = 0; co < Cout; co++)
            for (int ci = 0; ci < Cin; ci++)
            {
                __nv_bfloat16 b = __float2bfloat16_rn(hwio[(static_cast<size_t>(t) * Cin + ci) * Cout + co]);
                packed[(static_cast<size_t>(t) * N + co) * Cin + ci] = *reinterpret_cast<uint16_t*>(&b);
            }
    return packed;
}

ConvTc* ConvTcCreate(void const* d_in, void const* d_w_packed, int H, int W, int Cin, int Cout, bool out_bf16,
                     Act act, float alpha, int sm_count)
{
    auto op = std::make_unique<ConvTc>();
    TcParams& p = op->p;
    p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout;
    p.N = ConvTcPaddedCout(Cout);
    if (out_bf16 && p.N != Cout) throw std::runtime_error("bf16 output needs Cout to be a multiple of 16");
    p.KC = Cin < 64 ? Cin : 64;
    p.n_chunks = Cin / p.KC;
    p.layout_type = p.KC == 64 ? 2 : p.KC == 32 ? 4 : 6;
    p.sbo = 8 * p.KC * 2;
    p.tiles_per_row = (W + kTileM - 1) / kTileM;
    p.n_tiles = p.tiles_per_row * H;
    int cols = 32;
    while (cols < 2 * p.N) cols *= 2;
    p.tmem_cols = cols;
    p.out_bf16 = out_bf16 ? 1 : 0;
    p.act = act;
    p.alpha = alpha;
    p.debug = std::getenv("MLB200_TC_DEBUG") ? std::atoi(std::getenv("MLB200_TC_DEBUG")) : 0;
    p.halo = HaloMode(p.KC) ? 1 : 0;
    p.stage_bytes = static_cast<int>(StageBytes(p.KC));
    p.tap_stride = p.halo ? 128 : kTileM * p.KC * 2;
    p.stage_tx = p.halo ? (kTileM + 2) * 128 : 3 * kTileM * p.KC * 2;
    p.base_offset_mode = std::getenv("MLB200_TC_BASEOFF") ? std::atoi(std::getenv("MLB200_TC_BASEOFF")) : 0;
    p.stages = kMaxStages;
    while (p.stages > 2 && SmemBytes(p.N, Cin, p.KC, p.stages) > 226u * 1024u) p.stages--;
    op->smem = SmemBytes(p.N, Cin, p.KC, p.stages);
    if (op->smem > 226u * 1024u) throw std::runtime_error("filter does not fit in shared memory");
    op->grid = p.n_tiles < sm_count ? p.n_tiles : sm_count;

    // activation: [H][W][Cin] bf16, box = KC channels x 128 pixels x 1 row
    {
        cuuint64_t dims[3] = {static_cast<cuuint64_t>(Cin), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H)};
        cuuint64_t strides[2] = {static_cast<cuuint64_t>(Cin) * 2, static_cast<cuuint64_t>(W) * Cin * 2};
        cuuint32_t box[3] = {static_cast<cuuint32_t>(p.KC), static_cast<cuuint32_t>(p.halo ? kTileM + 2 : kTileM), 1};
        Encode(&op->in_map, d_in, 3, dims, strides, box, SwizzleFor(p.KC));
    }
    // filter: [9 * N][Cin] bf16, box = KC channels x N rows
    {
        cuuint64_t dims[2] = {static_cast<cuuint64_t>(Cin), static_cast<cuuint64_t>(9) * p.N};
        cuuint64_t strides[1] = {static_cast<cuuint64_t>(Cin) * 2};
        cuuint32_t box[2] = {static_cast<cuuint32_t>(p.KC), static_cast<cuuint32_t>(p.N)};
        Encode(&op->w_map, d_w_packed, 2, dims, strides, box, SwizzleFor(p.KC));
    }
    // per device: the attribute belongs to the function in the current context
    ML_CUDA_CHECK(cudaFuncSetAttribute(conv3x3_tc_kernel, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024));
    return op.release();
}

void ConvTcRun(ConvTc const* op, float const* bias, void* out, cudaStream_t stream)
{
    conv3x3_tc_kernel<<<op->grid, kTcThreads, op->smem, stream>>>(op->in_map, op->w_map, bias, out, op->p);
    ML_CUDA_CHECK(cudaGetLastError());
}

void ConvTcDestroy(ConvTc* op) { delete op; }

} // namespace nn
} // namespace ML
