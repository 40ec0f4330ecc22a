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
// Persistent kernel, one CTA per SM, 10 warps:
//   warp 0 (one lane)   TMA producer: the filter (once, stays resident in smem),
//                       then a ring of A stages, one per (tile, tap, Cin chunk)
//   warp 1 (one lane)   MMA issuer: Cin/16 tcgen05.mma per stage, tcgen05.commit
//                       frees the stage / publishes the accumulator
//   warps 2-9           epilogue: tcgen05.ld 32 lanes x 16 columns at a time,
//                       + bias, activation, pack to bf16 (or fp32 for the last
//                       layer), 16-byte stores; the accumulator is double
//                       buffered in TMEM so tile i's epilogue overlaps tile
//                       i+1's MMAs.
#include "nn_kernels.h"

#include "common.h"

#include <cuda.h>
#include <cudaTypedefs.h>
#include <cuda_bf16.h>

#include <cstdio>
#include <cstdlib>
#include <memory>
#include <stdexcept>
#include <string>
#include <vector>

namespace ML {
namespace nn {
namespace {

constexpr int kTileM = 128;          // output pixels per tile = TMEM lanes
constexpr int kEpiWarps = 8;         // two per TMEM lane quarter, each takes half of the columns
constexpr int kTcThreads = 64 + 32 * kEpiWarps;   // TMA warp + MMA warp + epilogue warps
constexpr int kMaxStages = 8;
constexpr int kMaxAcc = 4;
constexpr int kMaxChunks = 16;       // K chunks per filter row (tile-per-stage kernel)

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
    int n_acc;           // accumulators kept in TMEM (2 or 4): tile i's epilogue overlaps later tiles' MMAs
    int tmem_cols;       // allocated TMEM columns (power of two >= 32) = n_acc * N rounded up
    int out_bf16;        // 1: bf16 NHWC with N == Cout channels; 0: fp32 NHWC with Cout channels
    Act act;
    float alpha;
    int halo;            // 1: a stage is ONE box of 130 pixels (row y+kh-1, x0-1..x0+128) and the three
                         //    kw taps are row-shifted views of it (KC == 64 only); 0: three 128-pixel boxes
    int rps;             // filter rows per A stage: 3 (whole tile in one stage: one wait/fence/commit per
                         //    36 MMAs) when two such stages fit next to the filter, else 1
    int row_bytes;       // bytes of one filter row inside a stage, multiple of 1024
    int stage_bytes;     // bytes of one A stage = rps * row_bytes
    int tap_stride;      // bytes from tap kw to tap kw+1 inside a stage: 128 (halo) or 128*KC*2
    int stage_tx;        // bytes TMA delivers per stage
    unsigned long long* prof; // MLB200_TC_PROF=1: 8 cycle counters summed over CTAs (see ConvTcRun)
    int wait_ns;         // suspend-time hint of the mbarrier waits in the rows kernel (0: none)
    int rows_mode;       // 1: conv3x3_tc_rows_kernel (filter rows stacked along N), 0: conv3x3_tc_kernel
    int debug;           // MLB200_TC_DEBUG bit 0: no TMA loads of A, bit 1: no MMAs, bit 2: no epilogue math,
                         // bit 3: no TMA stores (timing experiments only; results are garbage)
    // Tile-per-stage kernel: K chunk c of a filter row multiplies the activation box at channel
    // a_coord[c] with filter chunk b_chunk[c] (of n_bchunks per tap).  Plain bf16: the identity.
    // Split mode (bf16x3, fp32-grade): activations and filter are stored as [hi | lo] bf16 halves
    // (x = hi + lo to 16 mantissa bits) and the chunk list runs hi*hi, lo*hi, hi*lo into the same
    // fp32 accumulator (the lo*lo term, 2^-16 relative, is dropped).
    int split;
    int out_split;       // 1: the fp32-output epilogue writes bf16 [hi: N | lo: N] halves per pixel instead
    int split_stage;     // out_split through the staging tiles + TMA store (when shared memory has room for them)
    int n_bchunks;
    int a_coord[kMaxChunks];
    int b_chunk[kMaxChunks];
    // Split mode, "wide" form (2N <= 256): the filter tile of a (tap, chunk) stacks the hi rows and the
    // lo rows along N, so ONE M128 x 2N MMA computes A_hi*B_hi into columns [0,N) and A_hi*B_lo into
    // [N,2N) while reading A_hi from shared memory once; A_lo*B_hi is an N-wide MMA on the tile's first
    // N rows.  The epilogue adds the two column halves.  14 KB of operand reads per K step instead of
    // 18 KB -- the tile-per-stage kernel is bound by exactly that.
    int wide;
    int acc_cols;        // TMEM columns per accumulator: 2N (wide) or N
    int chunk_wide[kMaxChunks];
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
// One probe of the phase without suspending: true when it has completed.
__device__ __forceinline__ bool bar_test(uint64_t* bar, uint32_t parity)
{
    uint32_t done;
    asm volatile(
        "{\n"
        ".reg .pred p;\n"
        "mbarrier.test_wait.parity.shared::cta.b64 p, [%1], %2;\n"
        "selp.u32 %0, 1, 0, p;\n"
        "}\n"
        : "=r"(done)
        : "r"(smem_addr(bar)), "r"(parity)
        : "memory");
    return done != 0;
}
// The same with a suspend-time hint (ns): the thread may stay suspended in hardware that long before
// try_wait returns false, so a waiting warp polls less often (MLB200_TC_WAIT_NS; 0 = no hint).
__device__ __forceinline__ void bar_wait(uint64_t* bar, uint32_t parity, uint32_t hint_ns)
{
    if (hint_ns == 0u) { bar_wait(bar, parity); return; }
    uint32_t done;
    do
    {
        asm volatile(
            "{\n"
            ".reg .pred p;\n"
            "mbarrier.try_wait.parity.shared::cta.b64 p, [%1], %2, %3;\n"
            "selp.u32 %0, 1, 0, p;\n"
            "}\n"
            : "=r"(done)
            : "r"(smem_addr(bar)), "r"(parity), "r"(hint_ns)
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
__device__ __forceinline__ void tma_store_3d(CUtensorMap const* map, void const* src, int c0, int c1, int c2)
{
    asm volatile(
        "cp.async.bulk.tensor.3d.global.shared::cta.tile.bulk_group [%0, {%2, %3, %4}], [%1];"
        ::"l"(map), "r"(smem_addr(src)), "r"(c0), "r"(c1), "r"(c2)
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
// One lane of a fully active warp (the lowest): the role loops below run with ALL lanes of their
// warp in uniform control flow and only the asynchronous instructions sit under this predicate, so
// the compiler keeps descriptors and addresses in uniform registers (a loop under `if (lane == 0)`
// costs an ELECT/R2UR/branch sequence per UTCHMMA and makes the issuing thread the bottleneck).
__device__ __forceinline__ bool elect_one()
{
    uint32_t pred;
    asm volatile(
        "{\n"
        ".reg .b32 rx;\n"
        ".reg .pred px;\n"
        "elect.sync rx|px, 0xffffffff;\n"
        "selp.u32 %0, 1, 0, px;\n"
        "}\n"
        : "=r"(pred));
    return pred != 0;
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
// D[tmem] (+)= A[smem] * B[smem], bf16 x bf16 -> fp32, issued by ONE thread for the CTA.  The
// descriptors are given as 32-bit halves: both share the constant high word, the low words are
// base + compile-time offset, so the single issuing warp spends ~3 uniform-datapath instructions per
// MMA (8 with 64-bit descriptor arithmetic and a run-time predicate -- that bounded the issue rate
// at N = 64).  ALWAYS_ACCUMULATE = false: D is overwritten when `accumulate` is 0 (a tile's first MMA).
template <bool ALWAYS_ACCUMULATE>
__device__ __forceinline__ void tc_mma_bf16_lohi(uint32_t d_tmem, uint32_t a_lo, uint32_t b_lo, uint32_t hi,
                                                 uint32_t idesc, uint32_t accumulate)
{
    if (ALWAYS_ACCUMULATE)
    {
        asm volatile(
            "{\n"
            ".reg .b64 da, db;\n"
            ".reg .pred p;\n"
            "setp.eq.b32 p, 0, 0;\n"
            "mov.b64 da, {%1, %3};\n"
            "mov.b64 db, {%2, %3};\n"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
            "}\n"
            ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc)
            : "memory");
    }
    else
    {
        asm volatile(
            "{\n"
            ".reg .b64 da, db;\n"
            ".reg .pred p;\n"
            "setp.ne.b32 p, %5, 0;\n"
            "mov.b64 da, {%1, %3};\n"
            "mov.b64 db, {%2, %3};\n"
            "tcgen05.mma.cta_group::1.kind::f16 [%0], da, db, %4, p;\n"
            "}\n"
            ::"r"(d_tmem), "r"(a_lo), "r"(b_lo), "r"(hi), "r"(idesc), "r"(accumulate)
            : "memory");
    }
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

// The transcendental activations stay out of line: inlined 16x per column group they made the
// epilogue loop body ~70 KB of SASS and the common ReLU path paid for it in instruction-cache misses.
__device__ __noinline__ float tc_act(float v, Act act, float alpha)
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
// 16-byte store to a shared-memory address (the compiler loses the address space of pointers derived
// from the dynamic shared-memory base and would emit generic ST.E.128)
__device__ __forceinline__ void sts128(uint32_t saddr, uint32_t a, uint32_t b, uint32_t c, uint32_t d)
{
    asm volatile("st.shared.v4.b32 [%0], {%1, %2, %3, %4};" ::"r"(saddr), "r"(a), "r"(b), "r"(c), "r"(d) : "memory");
}
__device__ __forceinline__ uint32_t pack_bf16(float lo, float hi)
{
    __nv_bfloat162 v = __floats2bfloat162_rn(lo, hi);
    return *reinterpret_cast<uint32_t*>(&v);
}

// MLB200_TC_PROF instrumentation: cycles a role spends in a region, summed over CTAs.
#define TC_PROF_T0() long long t0__ = P.prof ? clock64() : 0
#define TC_PROF_ADD(slot) do { if (P.prof) { long long t1__ = clock64(); prof_acc[slot] += t1__ - t0__; t0__ = t1__; } } while (0)

// ---------------------------------------------------------------- the kernel
template <int KC>
__global__ void __launch_bounds__(kTcThreads, 1)
conv3x3_tc_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap w_map,
                  const __grid_constant__ CUtensorMap out_map, float const* __restrict__ bias,
                  void* __restrict__ out, TcParams const P)
{
    extern __shared__ uint8_t smem_raw[];
    // 1024-byte alignment: the 128-byte swizzle pattern repeats every 1024 bytes
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~uintptr_t(1023u));
    uint32_t const b_rows = static_cast<uint32_t>(P.wide ? 2 * P.N : P.N);   // filter rows per (tap, chunk) tile
    uint32_t const b_tile_bytes = b_rows * KC * 2u;
    uint32_t const a_stage_bytes = static_cast<uint32_t>(P.stage_bytes);
    uint32_t const n_b_tiles = 9u * static_cast<uint32_t>(P.n_bchunks);
    // staged output (TMA store): bf16 layers write N channels per pixel, split layers 2N (hi | lo)
    bool const staged = P.out_bf16 || P.split_stage;
    uint32_t const NO = static_cast<uint32_t>(P.split_stage ? 2 * P.N : P.N);
    uint32_t const out_tile_bytes = staged ? static_cast<uint32_t>(kTileM) * NO * 2u : 0u;
    uint8_t* smem_b = smem;                                       // filter, resident
    uint8_t* smem_a = smem_b + n_b_tiles * b_tile_bytes;           // A ring
    uint8_t* smem_o = smem_a + static_cast<uint32_t>(P.stages) * a_stage_bytes;   // 2 output staging tiles
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_o + 2u * out_tile_bytes);
    uint64_t* full = bars;                       // [stages]   TMA -> MMA
    uint64_t* empty = bars + kMaxStages;         // [stages]   MMA -> TMA
    uint64_t* w_full = bars + 2 * kMaxStages;    // filter resident
    uint64_t* acc_full = w_full + 1;             // [n_acc]    MMA -> epilogue
    uint64_t* acc_empty = acc_full + kMaxAcc;    // [n_acc]    epilogue -> MMA
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(acc_empty + kMaxAcc);

    uint32_t const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;

    // bias (zero beyond Cout) staged once; the epilogue reads it as broadcast LDS.128
    __shared__ __align__(16) float s_bias[256];
    for (int c = threadIdx.x; c < P.N; c += kTcThreads) s_bias[c] = (bias != nullptr && c < P.Cout) ? bias[c] : 0.f;
    bool const relu_act = P.act == Act::Relu;
    bool const simple_act = P.act == Act::None || P.act == Act::Relu || P.act == Act::LeakyRelu;
    float const neg_slope = P.act == Act::None ? 1.f : (P.act == Act::LeakyRelu ? P.alpha : 0.f);

    if (warp == 0 && lane == 0)
    {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&in_map) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&w_map) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&out_map) : "memory");
        for (int s = 0; s < P.stages; s++) { bar_init(&full[s], 1); bar_init(&empty[s], 1); }
        bar_init(w_full, 1);
        for (int a = 0; a < P.n_acc; a++) { bar_init(&acc_full[a], 1); bar_init(&acc_empty[a], kEpiWarps); }
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
    long long prof_acc[5] = {0, 0, 0, 0, 0};

    // All three role loops run with every lane of their warp in uniform control flow; ring slot and
    // phase are carried incrementally (no division / modulo per stage) because each loop is a single
    // instruction stream whose length per tile bounds the kernel.
    if (warp == 0)
    {
        // ---- TMA producer
        if (elect_one())
        {
            bar_expect_tx(w_full, n_b_tiles * b_tile_bytes);
            for (uint32_t t = 0; t < 9u; t++)
                for (int ch = 0; ch < P.n_bchunks; ch++)
                    tma_load_2d(smem_b + (t * P.n_bchunks + ch) * b_tile_bytes, &w_map, ch * KC,
                                static_cast<int>(t * b_rows), w_full);
        }
        __syncwarp();
        uint32_t s = 0, ph = 0;
        int y = blockIdx.x / P.tiles_per_row, tx = blockIdx.x - y * P.tiles_per_row;
        int const dy = gridDim.x / P.tiles_per_row, dtx = gridDim.x - dy * P.tiles_per_row;
        TC_PROF_T0();
        for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x)
        {
            int const x0 = tx * kTileM;
            for (int kh0 = 0; kh0 < 3; kh0 += P.rps)
            {
                for (int ch = 0; ch < P.n_chunks; ch++)
                {
                    TC_PROF_ADD(1);
                    bar_wait(&empty[s], ph ^ 1u);
                    TC_PROF_ADD(0);
                    if (elect_one())
                    {
                        if (P.debug & 1)
                        {
                            bar_arrive(&full[s]);
                        }
                        else
                        {
                            bar_expect_tx(&full[s], static_cast<uint32_t>(P.stage_tx));
                            for (int r = 0; r < P.rps; r++)
                            {
                                // rows/pixels outside the image are zero-filled by TMA = SAME padding
                                uint8_t* dst = smem_a + s * a_stage_bytes + r * P.row_bytes;
                                int const iy = y + kh0 + r - 1;
                                if (P.halo)
                                {
                                    tma_load_3d(dst, &in_map, P.a_coord[ch], x0 - 1, iy, &full[s]);
                                }
                                else
                                {
                                    for (int kw = 0; kw < 3; kw++)
                                        tma_load_3d(dst + kw * P.tap_stride, &in_map, P.a_coord[ch], x0 + kw - 1, iy,
                                                    &full[s]);
                                }
                            }
                        }
                    }
                    __syncwarp();
                    if (++s == static_cast<uint32_t>(P.stages)) { s = 0; ph ^= 1u; }
                }
            }
            y += dy; tx += dtx;
            if (tx >= P.tiles_per_row) { tx -= P.tiles_per_row; y++; }
        }
        TC_PROF_ADD(1);
        if (P.prof && lane == 0)
        {
            atomicAdd(P.prof + 0, static_cast<unsigned long long>(prof_acc[0]));
            atomicAdd(P.prof + 1, static_cast<unsigned long long>(prof_acc[1]));
        }
    }
    else if (warp == 1)
    {
        // ---- MMA issuer.  Instruction descriptor (cute::UMMA::InstrDescriptor): D = fp32 [4,6),
        // A = B = bf16 [7,10) [10,13), both K-major, N>>3 at [17,23), M>>4 at [24,29).
        uint32_t const idesc_n = (1u << 4) | (1u << 7) | (1u << 10)
                               | (static_cast<uint32_t>(P.N >> 3) << 17)
                               | (static_cast<uint32_t>(kTileM >> 4) << 24);
        uint32_t const idesc_w = (1u << 4) | (1u << 7) | (1u << 10)
                               | (static_cast<uint32_t>((2 * P.N) >> 3) << 17)
                               | (static_cast<uint32_t>(kTileM >> 4) << 24);
        bar_wait(w_full, 0);
        tc_fence_after();
        // Everything of the two shared-memory descriptors except the start address is constant;
        // (addr >> 4) stays below 2^14 for any address inside the 227 KB window, so stages, taps and
        // K steps are plain additions to the low word.
        uint64_t const desc0 = umma_desc(0u, P.sbo, P.layout_type);
        uint32_t const desc_hi = static_cast<uint32_t>(desc0 >> 32);
        uint32_t const lo_flags = static_cast<uint32_t>(desc0);          // LBO field; start address added below
        uint32_t const a_lo0 = (smem_addr(smem_a) >> 4) + lo_flags, b_lo0 = (smem_addr(smem_b) >> 4) + lo_flags;
        uint32_t const stage_u = a_stage_bytes >> 4, tap_u = static_cast<uint32_t>(P.tap_stride) >> 4;
        uint32_t const b_tile_u = b_tile_bytes >> 4, row_u = static_cast<uint32_t>(P.row_bytes) >> 4;
        uint32_t const rps = static_cast<uint32_t>(P.rps), n_chunks = static_cast<uint32_t>(P.n_chunks);
        uint32_t const b_tap_u = static_cast<uint32_t>(P.n_bchunks) * b_tile_u, b_row_u = 3u * b_tap_u;
        uint32_t s = 0, ph = 0, acc = 0, aph = 0;
        TC_PROF_T0();
        for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x)
        {
            TC_PROF_ADD(2);
            bar_wait(&acc_empty[acc], aph ^ 1u);
            TC_PROF_ADD(0);
            tc_fence_after();
            uint32_t const d_tmem = tmem_base + acc * static_cast<uint32_t>(P.acc_cols);
            for (uint32_t kh0 = 0; kh0 < 3u; kh0 += rps)
            {
                for (uint32_t ch = 0; ch < n_chunks; ch++)
                {
                    TC_PROF_ADD(2);
                    bar_wait(&full[s], ph);
                    TC_PROF_ADD(1);
                    tc_fence_after();
                    if (elect_one())
                    {
                        if (!(P.debug & 2))
                        {
                            uint32_t const a_st = a_lo0 + s * stage_u;
                            uint32_t const b_st = b_lo0 + kh0 * b_row_u + static_cast<uint32_t>(P.b_chunk[ch]) * b_tile_u;
                            uint32_t const idesc = P.chunk_wide[ch] ? idesc_w : idesc_n;
#pragma unroll
                            for (uint32_t r = 0; r < 3u; r++)
                            {
                                if (r < rps)
                                {
#pragma unroll
                                    for (uint32_t kw = 0; kw < 3u; kw++)
                                    {
                                        // halo mode: tap kw = the same box read from pixel row kw on (start
                                        // kw * 128 bytes into the 1024-byte swizzle pattern).  Measured on
                                        // B200: the MMA applies the swizzle XOR to the ABSOLUTE shared-memory
                                        // address bits, as TMA does, so a row-shifted start needs no base
                                        // offset (descriptor bits [49,52) stay 0; setting them to
                                        // (start >> 7) & 7 gives wrong results).
#pragma unroll
                                        for (uint32_t k = 0; k < KC / 16; k++)
                                        {
                                            // +32 bytes along K inside the swizzle row = +2 in (addr >> 4)
                                            uint32_t const a_lo = a_st + r * row_u + kw * tap_u + 2u * k;
                                            uint32_t const b_lo = b_st + r * b_row_u + kw * b_tap_u + 2u * k;
                                            if (r == 0 && kw == 0 && k == 0)   // the tile's first MMA overwrites D
                                                tc_mma_bf16_lohi<false>(d_tmem, a_lo, b_lo, desc_hi, idesc, kh0 | ch);
                                            else
                                                tc_mma_bf16_lohi<true>(d_tmem, a_lo, b_lo, desc_hi, idesc, 1u);
                                        }
                                    }
                                }
                            }
                        }
                        TC_PROF_ADD(2);
                        tc_commit(&empty[s]);       // ONE commit per stage (12 or 36 MMAs at Cin = 64)
                        TC_PROF_ADD(3);
                    }
                    __syncwarp();
                    if (++s == static_cast<uint32_t>(P.stages)) { s = 0; ph ^= 1u; }
                    TC_PROF_ADD(4);
                }
            }
            if (elect_one()) tc_commit(&acc_full[acc]);   // accumulator complete
            __syncwarp();
            if (++acc == static_cast<uint32_t>(P.n_acc)) { acc = 0; aph ^= 1u; }
        }
        TC_PROF_ADD(4);
        if (P.prof && lane == 0)
        {
            for (int k = 0; k < 3; k++) atomicAdd(P.prof + 2 + k, static_cast<unsigned long long>(prof_acc[k]));
            atomicAdd(P.prof + 7, static_cast<unsigned long long>(prof_acc[3]));
            atomicAdd(P.prof + 8, static_cast<unsigned long long>(prof_acc[4]));
        }
    }
    else
    {
        // ---- epilogue warps 2..9: warp w may touch TMEM lanes 32*(w%4) .. +31 = tile rows (pixels).
        // bf16 output: each thread packs its pixel's channels and parks them in a swizzled staging tile
        // (conflict-free STS.128), then ONE thread hands the whole 128-pixel tile to TMA
        // (cp.async.bulk.tensor store: coalesced, asynchronous, clipped at the image edge).  Two staging
        // tiles alternate, so tile i's store drains while tile i+1 is being packed.
        uint32_t const q = warp & 3u;
        uint32_t const row = q * 32u + lane;
        bool const issuer = threadIdx.x == 64;               // warp 2, lane 0
        // the two warps of a lane quarter split the N/16 column units: [c_begin, c_end)
        int const units = P.N >> 4, half = static_cast<int>(warp - 2u) >> 2;
        int const c_begin = half == 0 ? 0 : ((units + 1) >> 1) << 4;
        int const c_end = half == 0 ? ((units + 1) >> 1) << 4 : P.N;
        int const OC = NO < 64u ? static_cast<int>(NO) : 64;  // channels per output box / staging chunk
        uint32_t const row_bytes = static_cast<uint32_t>(OC) * 2u;
        uint32_t const swz_mask = row_bytes == 128u ? 7u : (row_bytes == 64u ? 3u : 1u);
        uint32_t const chunk_bytes = static_cast<uint32_t>(kTileM) * row_bytes;
        uint32_t acc = 0, aph = 0, buf = 0;
        int y = blockIdx.x / P.tiles_per_row, tx = blockIdx.x - y * P.tiles_per_row;
        int const dy = gridDim.x / P.tiles_per_row, dtx = gridDim.x - dy * P.tiles_per_row;
        TC_PROF_T0();
        for (int tile = blockIdx.x; tile < P.n_tiles; tile += gridDim.x)
        {
            int const x0 = tx * kTileM;
            TC_PROF_ADD(1);
            bar_wait(&acc_full[acc], aph);
            TC_PROF_ADD(0);
            tc_fence_after();
            uint32_t const taddr = tmem_base + acc * static_cast<uint32_t>(P.acc_cols) + ((q * 32u) << 16);
            uint8_t* const stg = smem_o + buf * out_tile_bytes;
            int const x = x0 + static_cast<int>(row);
            size_t const pixel = static_cast<size_t>(y) * P.W + x;
            for (int c0 = c_begin; c0 < c_end && !(P.debug & 4); c0 += (P.wide ? 16 : 32))
            {
                // two 16-column loads in flight, one wait (N is a multiple of 16: the second
                // half is warp-uniformly skipped on the last odd chunk)
                uint32_t r[32];
                // wide (bf16x3): the second load is the same 16 channels' hi*lo half, N columns further on
                bool const two = P.wide || c0 + 16 < c_end;
                tc_ld16(taddr + static_cast<uint32_t>(c0), *reinterpret_cast<uint32_t(*)[16]>(&r[0]));
                if (two) tc_ld16(taddr + static_cast<uint32_t>(P.wide ? P.N + c0 : c0 + 16), *reinterpret_cast<uint32_t(*)[16]>(&r[16]));
                tc_ld_wait();
                if (P.wide)
                {
#pragma unroll
                    for (int j = 0; j < 16; j++) r[j] = __float_as_uint(__uint_as_float(r[j]) + __uint_as_float(r[16 + j]));
                }
#pragma unroll
                for (int h = 0; h < 2; h++)
                {
                    if (h == 1 && (!two || P.wide)) break;
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
                    if (relu_act)
                    {
#pragma unroll
                        for (int j = 0; j < 16; j++) v[j] = fmaxf(v[j], 0.f);
                    }
                    else if (simple_act)
                    {
                        // None / LeakyRelu without per-element branches: max(v,0) + slope * min(v,0)
#pragma unroll
                        for (int j = 0; j < 16; j++) v[j] = fmaf(neg_slope, fminf(v[j], 0.f), fmaxf(v[j], 0.f));
                    }
                    else
                    {
#pragma unroll
                        for (int j = 0; j < 16; j++) v[j] = tc_act(v[j], P.act, P.alpha);
                    }
                    // 16 packed channels starting at channel c = two 16-byte units of this pixel's row in staging
                    // chunk c / 64, under the swizzle TMA expects (address bits, 1024-aligned base)
                    auto stage16 = [&](uint32_t c, uint4 lo8, uint4 hi8) {
                        uint32_t const oc = c >> 6, unit = (c & 63u) >> 3;
                        uint32_t off0 = row * row_bytes + unit * 16u;
                        uint32_t off1 = off0 + 16u;
                        off0 ^= ((off0 >> 7) & swz_mask) << 4;
                        off1 ^= ((off1 >> 7) & swz_mask) << 4;
                        uint8_t* const base = stg + oc * chunk_bytes;
                        *reinterpret_cast<uint4*>(base + off0) = lo8;
                        *reinterpret_cast<uint4*>(base + off1) = hi8;
                    };
                    if (P.out_bf16)
                    {
                        stage16(static_cast<uint32_t>(cb),
                                make_uint4(pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                                           pack_bf16(v[6], v[7])),
                                make_uint4(pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]),
                                           pack_bf16(v[14], v[15])));
                    }
                    else if (P.split_stage)
                    {
                        // bf16x3 through the staging tile: hi at channel cb, lo at channel N + cb of the 2N-channel row
                        uint32_t hi[8], lo[8];
#pragma unroll
                        for (int j = 0; j < 8; j++)
                        {
                            __nv_bfloat162 hh = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                            float2 const hf = __bfloat1622float2(hh);
                            __nv_bfloat162 ll = __floats2bfloat162_rn(v[2 * j] - hf.x, v[2 * j + 1] - hf.y);
                            hi[j] = *reinterpret_cast<uint32_t*>(&hh);
                            lo[j] = *reinterpret_cast<uint32_t*>(&ll);
                        }
                        stage16(static_cast<uint32_t>(cb), make_uint4(hi[0], hi[1], hi[2], hi[3]),
                                make_uint4(hi[4], hi[5], hi[6], hi[7]));
                        stage16(static_cast<uint32_t>(P.N + cb), make_uint4(lo[0], lo[1], lo[2], lo[3]),
                                make_uint4(lo[4], lo[5], lo[6], lo[7]));
                    }
                    else if (P.out_split)
                    {
                        // bf16x3: v = hi + lo, hi to channels [cb, cb+16), lo to [N + cb, ...) of the pixel's row of
                        // 2N bf16: four 16-byte stores, whole 32-byte sectors
                        if (x < P.W)
                        {
                            uint32_t hi[8], lo[8];
#pragma unroll
                            for (int j = 0; j < 8; j++)
                            {
                                __nv_bfloat162 h = __floats2bfloat162_rn(v[2 * j], v[2 * j + 1]);
                                float2 const hf = __bfloat1622float2(h);
                                __nv_bfloat162 l = __floats2bfloat162_rn(v[2 * j] - hf.x, v[2 * j + 1] - hf.y);
                                hi[j] = *reinterpret_cast<uint32_t*>(&h);
                                lo[j] = *reinterpret_cast<uint32_t*>(&l);
                            }
                            uint8_t* const o = static_cast<uint8_t*>(out) + pixel * (static_cast<size_t>(P.N) * 4u)
                                             + static_cast<size_t>(cb) * 2u;
                            *reinterpret_cast<uint4*>(o) = make_uint4(hi[0], hi[1], hi[2], hi[3]);
                            *reinterpret_cast<uint4*>(o + 16) = make_uint4(hi[4], hi[5], hi[6], hi[7]);
                            uint8_t* const ol = o + static_cast<size_t>(P.N) * 2u;
                            *reinterpret_cast<uint4*>(ol) = make_uint4(lo[0], lo[1], lo[2], lo[3]);
                            *reinterpret_cast<uint4*>(ol + 16) = make_uint4(lo[4], lo[5], lo[6], lo[7]);
                        }
                    }
                    else if (x < P.W)
                    {
                        float* o = static_cast<float*>(out) + pixel * P.Cout;
                        if ((P.Cout & 3) == 0)
                        {
                            // 16-byte stores (buffers come from cudaMalloc; pixel * Cout + cb is a multiple of 4)
#pragma unroll
                            for (int j4 = 0; j4 < 4; j4++)
                                if (cb + 4 * j4 < P.Cout)
                                    *reinterpret_cast<float4*>(o + cb + 4 * j4) =
                                        make_float4(v[4 * j4], v[4 * j4 + 1], v[4 * j4 + 2], v[4 * j4 + 3]);
                        }
                        else
                        {
#pragma unroll
                            for (int j = 0; j < 16; j++)
                                if (cb + j < P.Cout) o[cb + j] = v[j];
                        }
                    }
                }
            }
            TC_PROF_ADD(2);
            // TMEM accumulator is free as soon as it has been read
            tc_fence_before();
            __syncwarp();
            if (lane == 0) bar_arrive(&acc_empty[acc]);
            if (++acc == static_cast<uint32_t>(P.n_acc)) { acc = 0; aph ^= 1u; }

            if (staged)
            {
                // generic-proxy writes -> visible to the async proxy (TMA) ; the issuer first makes sure
                // the store that last read the OTHER staging tile is done, so that after the barrier
                // every thread may start overwriting it for the next tile
                asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                TC_PROF_ADD(1);
                if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                TC_PROF_ADD(3);
                asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
                TC_PROF_ADD(4);
                if (issuer)
                {
                    for (int oc = 0; oc * 64 < static_cast<int>(NO); oc++)
                        tma_store_3d(&out_map, stg + oc * chunk_bytes, oc * 64, x0, y);
                    asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                }
                buf ^= 1u;
            }
            y += dy; tx += dtx;
            if (tx >= P.tiles_per_row) { tx -= P.tiles_per_row; y++; }
        }
        if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        TC_PROF_ADD(1);
        if (P.prof && warp == 2 && lane == 0)
        {
            atomicAdd(P.prof + 5, static_cast<unsigned long long>(prof_acc[0]));
            atomicAdd(P.prof + 6, static_cast<unsigned long long>(prof_acc[1]));
            for (int k = 2; k < 5; k++) atomicAdd(P.prof + 7 + k, static_cast<unsigned long long>(prof_acc[k]));
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

// ---------------------------------------------------------------- the row-stacked kernel
// Same convolution, organised around INPUT rows.  Input row r contributes to the three output rows
// r+1, r, r-1 through the filter rows kh = 0, 1, 2, so ONE MMA with the three filter rows stacked
// along N (N3 = 3 * N columns of TMEM = three output-row accumulators side by side) does the work
// of three MMAs of the tile-per-stage kernel above while reading the A operand from shared memory
// once instead of three times: 10 KB of operand reads per 128 x 192 x 16 MMA instead of 18 KB.
// The kernel above is bound by exactly that shared-memory bandwidth.
//
// A CTA walks DOWN a strip of 128-pixel tiles, one TMA box (one input row) per stage.  The
// accumulator of output row y lives in TMEM slot y mod 3; for input row r the stacked B operand
// must present filter rows (kh for slot 0, 1, 2) = a cyclic rotation of (1, 0, 2) that depends on
// r mod 3, so each kw tap keeps FIVE filter blocks [W1 W0 W2 W1 W0] in shared memory and the
// descriptor start picks the window.  Output row r+1 starts at input row r: the first K step of a
// row is issued as three N-wide MMAs whose accumulate flags differ per slot, the rest N3-wide.
// Output row r-1 is complete after input row r and is drained by the epilogue warps while the MMA
// pipe works on the CTA's second, independent strip ("pipe" 0 / 1, TMEM columns [0,3N) / [3N,6N)).
struct RowWalk
{
    int t, t_end, H;             // this pipe's share [t, t_end) of the tile sequence t = strip * H + y
    int strip, ya, yb;           // current segment: output rows [ya, yb) of one strip
    int r, r_last;               // current / last input row of the segment
    bool first;                  // r is the segment's first input row: every slot starts afresh

    __device__ __forceinline__ void Open()
    {
        if (t >= t_end) return;
        strip = t / H;
        ya = t - strip * H;
        yb = min(H, ya + (t_end - t));
        r = ya > 0 ? ya - 1 : 0;             // rows above / below the image contribute nothing (SAME padding)
        r_last = yb < H ? yb : H - 1;
        first = true;
    }
    __device__ __forceinline__ void Init(int q, int n_pipes, int n_tiles, int H_)
    {
        H = H_;
        t = static_cast<int>(static_cast<long long>(n_tiles) * q / n_pipes);
        t_end = static_cast<int>(static_cast<long long>(n_tiles) * (q + 1) / n_pipes);
        Open();
    }
    __device__ __forceinline__ bool Valid() const { return t < t_end; }
    __device__ __forceinline__ void Next()
    {
        first = false;
        if (++r > r_last) { t += yb - ya; Open(); }
    }
};

// EPI selects the epilogue at compile time (one straight-line loop body per variant keeps it inside the
// instruction cache): bit 0 = bf16 output through the TMA store (else fp32 direct stores), bits 1-2 =
// activation class 0 ReLU, 1 none / LeakyReLU, 2 anything else (out of line).
template <int KC, int EPI>
__global__ void __launch_bounds__(kTcThreads, 1)
conv3x3_tc_rows_kernel(const __grid_constant__ CUtensorMap in_map, const __grid_constant__ CUtensorMap w_map,
                       const __grid_constant__ CUtensorMap out_map, float const* __restrict__ bias,
                       void* __restrict__ out, TcParams const P)
{
    constexpr bool kOutBf16 = (EPI & 1) != 0;
    constexpr int kActClass = EPI >> 1;
    extern __shared__ uint8_t smem_raw[];
    uint8_t* smem = reinterpret_cast<uint8_t*>((reinterpret_cast<uintptr_t>(smem_raw) + 1023u) & ~uintptr_t(1023u));
    uint32_t const b_blk_bytes = static_cast<uint32_t>(P.N) * KC * 2u;      // one filter row of one kw tap
    uint32_t const a_stage_bytes = static_cast<uint32_t>(P.stage_bytes);
    uint32_t const out_tile_bytes = kOutBf16 ? static_cast<uint32_t>(kTileM) * P.N * 2u : 0u;
    uint8_t* smem_b = smem;                                       // [kw][W1 W0 W2 W1 W0] filter blocks, resident
    uint8_t* smem_a = smem_b + 15u * b_blk_bytes;                 // ring of input rows
    uint8_t* smem_o = smem_a + static_cast<uint32_t>(P.stages) * a_stage_bytes;   // 2 output staging tiles
    uint64_t* bars = reinterpret_cast<uint64_t*>(smem_o + 2u * out_tile_bytes);
    uint64_t* full = bars;                       // [stages]   TMA -> MMA
    uint64_t* done = bars + kMaxStages;          // [stages]   MMA -> TMA and MMA -> epilogue: the stage's input row has been
                                                 //            accumulated (ONE tcgen05.commit per row: the stage is free for
                                                 //            the producer AND the rows it completed can be drained)
    uint64_t* w_full = bars + 2 * kMaxStages;    // filter resident
    uint64_t* slot_free = w_full + 1;            // [2 pipes]  epilogue -> MMA: the rows it completed are drained
    uint32_t* tmem_slot = reinterpret_cast<uint32_t*>(slot_free + 2 * kMaxAcc);

    uint32_t const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    __shared__ __align__(16) float s_bias[256];
    for (int c = threadIdx.x; c < P.N; c += kTcThreads) s_bias[c] = (bias != nullptr && c < P.Cout) ? bias[c] : 0.f;
    float const neg_slope = P.act == Act::None ? 1.f : (P.act == Act::LeakyRelu ? P.alpha : 0.f);

    if (warp == 0 && lane == 0)
    {
        asm volatile("prefetch.tensormap [%0];" ::"l"(&in_map) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&w_map) : "memory");
        asm volatile("prefetch.tensormap [%0];" ::"l"(&out_map) : "memory");
        for (int s = 0; s < P.stages; s++) { bar_init(&full[s], 1); bar_init(&done[s], 1); }
        bar_init(w_full, 1);
        for (int a = 0; a < 2; a++) bar_init(&slot_free[a], kEpiWarps);
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
    long long prof_acc[8] = {0, 0, 0, 0, 0, 0, 0, 0};

    // Every role steps identical copies of the two pipes' walks, alternating pipe 0 / pipe 1.
    RowWalk walk[2];
    walk[0].Init(2 * blockIdx.x, 2 * gridDim.x, P.n_tiles, P.H);
    walk[1].Init(2 * blockIdx.x + 1, 2 * gridDim.x, P.n_tiles, P.H);
    uint32_t const N = static_cast<uint32_t>(P.N);
    uint32_t const wait_ns = static_cast<uint32_t>(P.wait_ns);

    if (warp == 0)
    {
        // ---- TMA producer: the filter blocks once, then one input row per stage
        if (elect_one())
        {
            bar_expect_tx(w_full, 15u * b_blk_bytes);
            for (int kw = 0; kw < 3; kw++)
                for (int j = 0; j < 5; j++)
                {
                    int const kh = j == 0 || j == 3 ? 1 : (j == 1 || j == 4 ? 0 : 2);
                    tma_load_2d(smem_b + (kw * 5 + j) * b_blk_bytes, &w_map, 0, (kh * 3 + kw) * P.N, w_full);
                }
        }
        __syncwarp();
        uint32_t s = 0, ph = 0;
        TC_PROF_T0();
        while (walk[0].Valid() || walk[1].Valid())
        {
#pragma unroll
            for (int p = 0; p < 2; p++)
            {
                if (!walk[p].Valid()) continue;
                TC_PROF_ADD(1);
                bar_wait(&done[s], ph ^ 1u, wait_ns);
                TC_PROF_ADD(0);
                if (elect_one())
                {
                    if (P.debug & 1)
                    {
                        bar_arrive(&full[s]);
                    }
                    else
                    {
                        bar_expect_tx(&full[s], static_cast<uint32_t>(P.stage_tx));
                        uint8_t* dst = smem_a + s * a_stage_bytes;
                        int const x0 = walk[p].strip * kTileM, iy = walk[p].r;
                        if (P.halo)
                        {
                            tma_load_3d(dst, &in_map, 0, x0 - 1, iy, &full[s]);   // pixels outside the row: zeros
                        }
                        else
                        {
                            for (int kw = 0; kw < 3; kw++)
                                tma_load_3d(dst + kw * P.tap_stride, &in_map, 0, x0 + kw - 1, iy, &full[s]);
                        }
                    }
                }
                __syncwarp();
                if (++s == static_cast<uint32_t>(P.stages)) { s = 0; ph ^= 1u; }
                walk[p].Next();
            }
        }
        TC_PROF_ADD(1);
        if (P.prof && lane == 0)
        {
            atomicAdd(P.prof + 0, static_cast<unsigned long long>(prof_acc[0]));
            atomicAdd(P.prof + 1, static_cast<unsigned long long>(prof_acc[1]));
        }
    }
    else if (warp == 1)
    {
        // ---- MMA issuer
        uint32_t const idesc_base = (1u << 4) | (1u << 7) | (1u << 10) | (static_cast<uint32_t>(kTileM >> 4) << 24);
        uint32_t const idesc1 = idesc_base | ((N >> 3) << 17);            // one slot
        uint32_t const idesc3 = idesc_base | (((3u * N) >> 3) << 17);     // the three slots of a pipe
        bar_wait(w_full, 0);
        tc_fence_after();
        uint64_t const desc0 = umma_desc(0u, P.sbo, P.layout_type);
        uint32_t const desc_hi = static_cast<uint32_t>(desc0 >> 32);
        uint32_t const lo_flags = static_cast<uint32_t>(desc0);
        uint32_t const a_lo0 = (smem_addr(smem_a) >> 4) + lo_flags, b_lo0 = (smem_addr(smem_b) >> 4) + lo_flags;
        uint32_t const stage_u = a_stage_bytes >> 4, tap_u = static_cast<uint32_t>(P.tap_stride) >> 4;
        uint32_t const b_blk_u = b_blk_bytes >> 4, b_kw_u = 5u * b_blk_u;
        uint32_t s = 0, ph = 0;
        uint32_t ev[2] = {0u, 0u};       // input rows issued per pipe (= barrier phases)
        // The waits of step k+1 (its stage landed, the slot it starts afresh drained) are taken in the MIDDLE of
        // step k's MMAs, while the tensor pipe still has the first ten queued: the MMA queue is short, and
        // every cycle this thread spends between the last MMA of one row and the first of the next (barrier
        // latency, fence, loop bookkeeping) is a cycle the pipe runs dry.
        bool pre_waited = false;
        TC_PROF_T0();
        while (walk[0].Valid() || walk[1].Valid())
        {
#pragma unroll
            for (int p = 0; p < 2; p++)
            {
                if (!walk[p].Valid()) continue;
                TC_PROF_ADD(2);
                if (!pre_waited)
                {
                    // the slot this row starts afresh held output row r-2, completed by the pipe's previous row
                    if (ev[p] > 0) bar_wait(&slot_free[p], (ev[p] - 1u) & 1u, wait_ns);
                    bar_wait(&full[s], ph, wait_ns);
                    tc_fence_after();
                }
                TC_PROF_ADD(0);
                uint32_t const c = static_cast<uint32_t>(walk[p].r) % 3u;
                uint32_t const win = c == 0u ? 0u : (c == 1u ? 2u : 1u);      // window start in [W1 W0 W2 W1 W0]
                uint32_t const fresh = c == 2u ? 0u : c + 1u;                  // slot of output row r+1
                uint32_t const all_fresh = walk[p].first ? 1u : 0u;
                uint32_t const a_st = a_lo0 + s * stage_u;
                uint32_t const b_st = b_lo0 + win * b_blk_u;
                uint32_t const d_tmem = tmem_base + static_cast<uint32_t>(p) * 3u * N;
                uint64_t* const done_bar = &done[s];
                bool const mma_on = !(P.debug & 2);
                if (elect_one() && mma_on)
                {
#pragma unroll
                    for (uint32_t sl = 0; sl < 3u; sl++)
                        tc_mma_bf16_lohi<false>(d_tmem + sl * N, a_st, b_st + sl * b_blk_u, desc_hi, idesc1,
                                                (all_fresh | (sl == fresh ? 1u : 0u)) ^ 1u);
#pragma unroll
                    for (uint32_t kw = 0; kw < 2u; kw++)
                    {
#pragma unroll
                        for (uint32_t k = 0; k < KC / 16; k++)
                        {
                            if (kw == 0 && k == 0) continue;
                            tc_mma_bf16_lohi<true>(d_tmem, a_st + kw * tap_u + 2u * k, b_st + kw * b_kw_u + 2u * k,
                                                   desc_hi, idesc3, 1u);
                        }
                    }
                }
                __syncwarp();
                TC_PROF_ADD(2);
                // bookkeeping and a probe of the next step's barriers, under cover of the queued MMAs.  The next
                // step is the other pipe's unless that one is exhausted; a step on the SAME pipe depends on this
                // row's own commit (its slot_free follows from it), so it is never probed early.
                if (++s == static_cast<uint32_t>(P.stages)) { s = 0; ph ^= 1u; }
                ev[p]++;
                walk[p].Next();
                pre_waited = false;
                if (walk[1 - p].Valid())
                {
                    // only if both have already happened: blocking here would hold this row's last MMAs back
                    bool ok = ev[1 - p] == 0 || bar_test(&slot_free[1 - p], (ev[1 - p] - 1u) & 1u);
                    ok = ok && bar_test(&full[s], ph);
                    if (ok)
                    {
                        tc_fence_after();
                        pre_waited = true;
                    }
                }
                TC_PROF_ADD(1);
                if (elect_one())
                {
                    if (mma_on)
                    {
#pragma unroll
                        for (uint32_t k = 0; k < KC / 16; k++)
                            tc_mma_bf16_lohi<true>(d_tmem, a_st + 2u * tap_u + 2u * k, b_st + 2u * b_kw_u + 2u * k,
                                                   desc_hi, idesc3, 1u);
                    }
                    tc_commit(done_bar);
                }
                __syncwarp();
                TC_PROF_ADD(3);
            }
        }
        TC_PROF_ADD(4);
        if (P.prof && lane == 0)
        {
            for (int k = 0; k < 3; k++) atomicAdd(P.prof + 2 + k, static_cast<unsigned long long>(prof_acc[k]));
            atomicAdd(P.prof + 7, static_cast<unsigned long long>(prof_acc[3]));
            atomicAdd(P.prof + 8, static_cast<unsigned long long>(prof_acc[4]));
        }
    }
    else
    {
        // ---- epilogue warps 2..9 (lane quarter q = warp % 4, two warps per quarter split the columns)
        uint32_t const q = warp & 3u;
        uint32_t const row = q * 32u + lane;
        bool const issuer = threadIdx.x == 64;
        int const units = P.N >> 4, half = static_cast<int>(warp - 2u) >> 2;
        int const c_begin = half == 0 ? 0 : ((units + 1) >> 1) << 4;
        int const c_end = half == 0 ? ((units + 1) >> 1) << 4 : P.N;
        int const OC = P.N < 64 ? P.N : 64;
        uint32_t const row_bytes = static_cast<uint32_t>(OC) * 2u;
        uint32_t const swz_mask = row_bytes == 128u ? 7u : (row_bytes == 64u ? 3u : 1u);
        uint32_t const chunk_bytes = static_cast<uint32_t>(kTileM) * row_bytes;
        uint32_t buf = 0;
        uint32_t s = 0, ph = 0;          // the stage ring, stepped exactly like the producer and the MMA warp do
        TC_PROF_T0();
        while (walk[0].Valid() || walk[1].Valid())
        {
#pragma unroll
            for (int p = 0; p < 2; p++)
            {
                if (!walk[p].Valid()) continue;
                // output rows this input row completes: r-1, and r itself on the image's last row
                int const r = walk[p].r;
                bool const above = r - 1 >= walk[p].ya;
                int const y_begin = above ? r - 1 : r;           // the completed rows are consecutive
                int const nd = (above ? 1 : 0) + ((r == P.H - 1 && r < walk[p].yb) ? 1 : 0);
                int const x0 = walk[p].strip * kTileM;
                int const x = x0 + static_cast<int>(row);
                TC_PROF_ADD(1);
                bar_wait(&done[s], ph, wait_ns);
                TC_PROF_ADD(0);
                tc_fence_after();
                for (int d = 0; d < nd; d++)
                {
                    int const y = y_begin + d;
                    uint32_t const slot = static_cast<uint32_t>(y) % 3u;
                    uint32_t const taddr = tmem_base + (static_cast<uint32_t>(p) * 3u + slot) * N + ((q * 32u) << 16);
                    uint8_t* const stg = smem_o + buf * out_tile_bytes;
                    size_t const pixel = static_cast<size_t>(y) * P.W + x;
                    for (int c0 = c_begin; c0 < c_end && !(P.debug & 4); c0 += 32)
                    {
                        uint32_t v32[32];
                        bool const two = c0 + 16 < c_end;
                        tc_ld16(taddr + static_cast<uint32_t>(c0), *reinterpret_cast<uint32_t(*)[16]>(&v32[0]));
                        if (two) tc_ld16(taddr + static_cast<uint32_t>(c0 + 16), *reinterpret_cast<uint32_t(*)[16]>(&v32[16]));
                        tc_ld_wait();
                        TC_PROF_ADD(5);
#pragma unroll
                        for (int h = 0; h < 2; h++)
                        {
                            if (h == 1 && !two) break;
                            int const cb = c0 + 16 * h;
                            float v[16];
#pragma unroll
                            for (int j4 = 0; j4 < 4; j4++)
                            {
                                float4 const b = *reinterpret_cast<float4 const*>(&s_bias[cb + 4 * j4]);   // broadcast LDS.128
                                v[4 * j4 + 0] = __uint_as_float(v32[16 * h + 4 * j4 + 0]) + b.x;
                                v[4 * j4 + 1] = __uint_as_float(v32[16 * h + 4 * j4 + 1]) + b.y;
                                v[4 * j4 + 2] = __uint_as_float(v32[16 * h + 4 * j4 + 2]) + b.z;
                                v[4 * j4 + 3] = __uint_as_float(v32[16 * h + 4 * j4 + 3]) + b.w;
                            }
                            if constexpr (kActClass == 0)
                            {
#pragma unroll
                                for (int j = 0; j < 16; j++) v[j] = fmaxf(v[j], 0.f);
                            }
                            else if constexpr (kActClass == 1)
                            {
#pragma unroll
                                for (int j = 0; j < 16; j++) v[j] = fmaf(neg_slope, fminf(v[j], 0.f), fmaxf(v[j], 0.f));
                            }
                            else
                            {
#pragma unroll
                                for (int j = 0; j < 16; j++) v[j] = tc_act(v[j], P.act, P.alpha);
                            }
                            if constexpr (kOutBf16)
                            {
                                uint32_t const oc = static_cast<uint32_t>(cb) >> 6;
                                uint32_t const unit = (static_cast<uint32_t>(cb) & 63u) >> 3;
                                uint32_t off0 = row * row_bytes + unit * 16u;
                                uint32_t off1 = off0 + 16u;
                                off0 ^= ((off0 >> 7) & swz_mask) << 4;
                                off1 ^= ((off1 >> 7) & swz_mask) << 4;
                                uint32_t const base = smem_addr(stg) + oc * chunk_bytes;
                                sts128(base + off0, pack_bf16(v[0], v[1]), pack_bf16(v[2], v[3]), pack_bf16(v[4], v[5]),
                                       pack_bf16(v[6], v[7]));
                                sts128(base + off1, pack_bf16(v[8], v[9]), pack_bf16(v[10], v[11]), pack_bf16(v[12], v[13]),
                                       pack_bf16(v[14], v[15]));
                            }
                            else
                            {
                              if (x < P.W)
                              {
                                float* o = static_cast<float*>(out) + pixel * P.Cout;
#pragma unroll
                                for (int j = 0; j < 16; j++)
                                    if (cb + j < P.Cout) o[cb + j] = v[j];
                              }
                            }
                        }
                    }
                    TC_PROF_ADD(2);
                    if (d == nd - 1)
                    {
                        // every completed row of this event has been read out of TMEM
                        tc_fence_before();
                        __syncwarp();
                        if (lane == 0) bar_arrive(&slot_free[p]);
                    }
                    TC_PROF_ADD(6);
                    if constexpr (kOutBf16)
                    {
                        asm volatile("fence.proxy.async.shared::cta;" ::: "memory");
                        TC_PROF_ADD(7);
                        if (issuer) asm volatile("cp.async.bulk.wait_group.read 0;" ::: "memory");
                        TC_PROF_ADD(3);
                        asm volatile("bar.sync 1, %0;" ::"n"(32 * kEpiWarps) : "memory");
                        TC_PROF_ADD(4);
                        if (issuer && !(P.debug & 8))
                        {
                            for (int oc = 0; oc * 64 < P.N; oc++)
                                tma_store_3d(&out_map, stg + oc * chunk_bytes, oc * 64, x0, y);
                            asm volatile("cp.async.bulk.commit_group;" ::: "memory");
                        }
                        buf ^= 1u;
                    }
                }
                if (nd == 0)
                {
                    tc_fence_before();
                    __syncwarp();
                    if (lane == 0) bar_arrive(&slot_free[p]);
                }
                if (++s == static_cast<uint32_t>(P.stages)) { s = 0; ph ^= 1u; }
                walk[p].Next();
            }
        }
        if (issuer) asm volatile("cp.async.bulk.wait_group 0;" ::: "memory");
        TC_PROF_ADD(1);
        if (P.prof && warp == 2 && lane == 0)
        {
            atomicAdd(P.prof + 5, static_cast<unsigned long long>(prof_acc[0]));
            atomicAdd(P.prof + 6, static_cast<unsigned long long>(prof_acc[1]));
            for (int k = 2; k < 8; k++) atomicAdd(P.prof + 7 + k, static_cast<unsigned long long>(prof_acc[k]));
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

int EnvInt(char const* name, int fallback)
{
    char const* v = std::getenv(name);
    return v != nullptr ? std::atoi(v) : fallback;
}

// Halo mode: one box of 130 pixels per input row, the kw taps are row-shifted views of it.
// Works under all three swizzles (128/64/32-byte rows).  MLB200_TC_HALO=0: three 128-pixel boxes instead.
bool HaloMode(int /*KC*/)
{
    return EnvInt("MLB200_TC_HALO", 1) != 0;
}

size_t StageBytes(int KC)
{
    // halo: 130 pixels x KC*2 B, rounded up to the 1024-byte swizzle pattern; else 3 boxes of 128 pixels
    size_t const halo = (static_cast<size_t>(kTileM + 2) * KC * 2 + 1023) / 1024 * 1024;
    return HaloMode(KC) ? halo : static_cast<size_t>(3) * kTileM * KC * 2;
}

// staged_channels: channels per pixel of the two output staging tiles (0: direct stores, no staging)
size_t SmemBytes(int N, int Cin, int KC, int stages, int staged_channels, int rps = 1)
{
    size_t b = static_cast<size_t>(9) * N * Cin * 2;
    size_t a = static_cast<size_t>(stages) * StageBytes(KC) * rps;
    size_t o = static_cast<size_t>(2) * kTileM * staged_channels * 2;    // two staging tiles for the TMA store
    return 1024 /*alignment slack*/ + b + a + o + (2 * kMaxStages + 1 + 2 * kMaxAcc) * sizeof(uint64_t) + 16;
}

size_t SmemBytesRows(int N, int KC, int stages, bool out_bf16)
{
    size_t b = static_cast<size_t>(15) * N * KC * 2;                    // 3 kw taps x [W1 W0 W2 W1 W0]
    size_t a = static_cast<size_t>(stages) * StageBytes(KC);            // one input row per stage
    size_t o = out_bf16 ? static_cast<size_t>(2) * kTileM * N * 2 : 0;
    return 1024 + b + a + o + (2 * kMaxStages + 1 + 2 * kMaxAcc) * sizeof(uint64_t) + 16;
}

// The row-stacked kernel needs the whole Cin in one swizzle row and three accumulators of a pipe
// inside one MMA's N range; MLB200_TC_ROWS=0 keeps the tile-per-stage kernel (for comparison).
bool RowsModeFits(int N, int Cin, bool out_bf16)
{
    if (EnvInt("MLB200_TC_ROWS", 1) == 0 || Cin > 64 || 3 * N > 256) return false;
    return SmemBytesRows(N, Cin, 3, out_bf16) <= 226u * 1024u;
}

using RowsKernel = void (*)(CUtensorMap, CUtensorMap, CUtensorMap, float const*, void*, TcParams);

template <int KC>
RowsKernel RowsKernelForKc(int epi)
{
    switch (epi)
    {
        case 0: return conv3x3_tc_rows_kernel<KC, 0>;
        case 1: return conv3x3_tc_rows_kernel<KC, 1>;
        case 2: return conv3x3_tc_rows_kernel<KC, 2>;
        case 3: return conv3x3_tc_rows_kernel<KC, 3>;
        case 4: return conv3x3_tc_rows_kernel<KC, 4>;
        default: return conv3x3_tc_rows_kernel<KC, 5>;
    }
}

RowsKernel RowsKernelFor(TcParams const& p)
{
    int const act_class = p.act == Act::Relu ? 0 : (p.act == Act::None || p.act == Act::LeakyRelu) ? 1 : 2;
    int const epi = (act_class << 1) | (p.out_bf16 ? 1 : 0);
    return p.KC == 64 ? RowsKernelForKc<64>(epi) : p.KC == 32 ? RowsKernelForKc<32>(epi) : RowsKernelForKc<16>(epi);
}

} // namespace

struct ConvTc
{
    CUtensorMap in_map;
    CUtensorMap w_map;
    CUtensorMap out_map;
    TcParams p;
    size_t smem;
    int grid;
};

int ConvTcPaddedCout(int Cout) { return (Cout + 15) / 16 * 16; }

bool ConvTcSupported(int KH, int KW, int stride, int dilation, bool same_padding, int Cin, int Cout, bool split)
{
    if (KH != 3 || KW != 3 || stride != 1 || dilation != 1 || !same_padding) return false;
    if (!(Cin == 16 || Cin == 32 || (Cin > 0 && Cin % 64 == 0))) return false;
    int const N = ConvTcPaddedCout(Cout);
    if (Cout < 1 || N > 256) return false;
    int const KC = Cin < 64 ? Cin : 64;
    if ((split ? 3 : 1) * (Cin / KC) > kMaxChunks) return false;
    if (split && N > 256) return false;
    // split: the resident filter holds the hi and the lo half, the output is fp32 (no staging tiles);
    // three ring stages at least, or the MMA pipe waits on every TMA round trip
    if (split) return SmemBytes(N, 2 * Cin, KC, 3, 0) <= 226u * 1024u;
    return SmemBytes(N, Cin, KC, 2, N) <= 226u * 1024u;
}

std::vector<uint16_t> ConvTcPackWeights(float const* hwio, int Cin, int Cout, int Cin_pad)
{
    int const N = ConvTcPaddedCout(Cout);
    std::vector<uint16_t> packed(static_cast<size_t>(9) * N * Cin_pad, 0);
    for (int t = 0; t < 9; t++)
        for (int co = 0; co < Cout; co++)
            for (int ci = 0; ci < Cin; ci++)
            {
                __nv_bfloat16 b = __float2bfloat16_rn(hwio[(static_cast<size_t>(t) * Cin + ci) * Cout + co]);
                packed[(static_cast<size_t>(t) * N + co) * Cin_pad + ci] = *reinterpret_cast<uint16_t*>(&b);
            }
    return packed;
}

bool ConvTcSplitWide(int Cout)
{
    return 2 * ConvTcPaddedCout(Cout) <= 256 && EnvInt("MLB200_X3_WIDE", 1) != 0;
}

std::vector<uint16_t> ConvTcPackWeightsSplit(float const* hwio, int Cin, int Cout, int Cin_pad)
{
    int const N = ConvTcPaddedCout(Cout);
    std::vector<uint16_t> packed(static_cast<size_t>(9) * N * 2 * Cin_pad, 0);
    if (ConvTcSplitWide(Cout))
    {
        // wide form: [tap][hi rows: N | lo rows: N][Cin_pad]
        for (int t = 0; t < 9; t++)
            for (int co = 0; co < Cout; co++)
                for (int ci = 0; ci < Cin; ci++)
                {
                    float const w = hwio[(static_cast<size_t>(t) * Cin + ci) * Cout + co];
                    __nv_bfloat16 hi = __float2bfloat16_rn(w);
                    __nv_bfloat16 lo = __float2bfloat16_rn(w - __bfloat162float(hi));
                    packed[(static_cast<size_t>(t) * 2 * N + co) * Cin_pad + ci] = *reinterpret_cast<uint16_t*>(&hi);
                    packed[(static_cast<size_t>(t) * 2 * N + N + co) * Cin_pad + ci] = *reinterpret_cast<uint16_t*>(&lo);
                }
        return packed;
    }
    // [tap][co][hi: Cin_pad | lo: Cin_pad], w = hi + lo up to 2^-17 relative
    for (int t = 0; t < 9; t++)
        for (int co = 0; co < Cout; co++)
            for (int ci = 0; ci < Cin; ci++)
            {
                float const w = hwio[(static_cast<size_t>(t) * Cin + ci) * Cout + co];
                __nv_bfloat16 hi = __float2bfloat16_rn(w);
                __nv_bfloat16 lo = __float2bfloat16_rn(w - __bfloat162float(hi));
                size_t const row = (static_cast<size_t>(t) * N + co) * 2 * Cin_pad;
                packed[row + ci] = *reinterpret_cast<uint16_t*>(&hi);
                packed[row + Cin_pad + ci] = *reinterpret_cast<uint16_t*>(&lo);
            }
    return packed;
}

ConvTc* ConvTcCreate(void const* d_in, void const* d_w_packed, void* d_out_bf16, int H, int W, int Cin, int Cout,
                     bool out_bf16, Act act, float alpha, int sm_count, bool split, bool out_split)
{
    auto op = std::make_unique<ConvTc>();
    TcParams& p = op->p;
    p.H = H; p.W = W; p.Cin = Cin; p.Cout = Cout;
    p.N = ConvTcPaddedCout(Cout);
    if (out_bf16 && p.N != Cout) throw std::runtime_error("bf16 output needs Cout to be a multiple of 16");
    if (split && out_bf16) throw std::runtime_error("split (bf16x3) layers write fp32");
    p.KC = Cin < 64 ? Cin : 64;
    int const n_real = Cin / p.KC;
    int const Ct = split ? 2 * Cin : Cin;      // channels of the activation and filter tensors ([hi | lo] when split)
    p.split = split ? 1 : 0;
    p.out_split = out_split ? 1 : 0;
    if (out_split && (out_bf16 || p.N != Cout)) throw std::runtime_error("split output needs fp32 mode and Cout % 16 == 0");
    p.wide = (split && ConvTcSplitWide(Cout)) ? 1 : 0;
    p.acc_cols = p.wide ? 2 * p.N : p.N;
    p.n_chunks = p.wide ? 2 * n_real : split ? 3 * n_real : n_real;
    p.n_bchunks = (split && !p.wide) ? 2 * n_real : n_real;
    if (p.n_chunks > kMaxChunks) throw std::runtime_error("too many K chunks");
    for (int c = 0; c < kMaxChunks; c++) { p.a_coord[c] = 0; p.b_chunk[c] = 0; p.chunk_wide[c] = 0; }
    for (int c = 0; c < p.n_chunks; c++)
    {
        if (p.wide)
        {
            // per real chunk: A_hi x [B_hi | B_lo] (2N wide, first: it overwrites every column), then A_lo x B_hi
            int const r = c >> 1, lo = c & 1;
            p.a_coord[c] = (lo ? Cin : 0) + r * p.KC;
            p.b_chunk[c] = r;
            p.chunk_wide[c] = lo ? 0 : 1;
        }
        else
        {
            int const term = c / n_real, r = c - term * n_real;   // split terms: hi*hi, lo*hi, hi*lo
            p.a_coord[c] = (term == 1 ? Cin : 0) + r * p.KC;
            p.b_chunk[c] = (term == 2 ? n_real : 0) + r;
        }
    }
    p.layout_type = p.KC == 64 ? 2 : p.KC == 32 ? 4 : 6;
    p.sbo = 8 * p.KC * 2;
    p.tiles_per_row = (W + kTileM - 1) / kTileM;
    p.n_tiles = p.tiles_per_row * H;
    int const want_acc = std::getenv("MLB200_TC_ACC") ? std::atoi(std::getenv("MLB200_TC_ACC")) : kMaxAcc;
    p.n_acc = (want_acc >= 4 && 4 * p.acc_cols <= 512) ? 4 : 2;
    int cols = 32;
    while (cols < p.n_acc * p.acc_cols) cols *= 2;
    p.tmem_cols = cols;
    p.out_bf16 = out_bf16 ? 1 : 0;
    p.act = act;
    p.alpha = alpha;
    p.debug = std::getenv("MLB200_TC_DEBUG") ? std::atoi(std::getenv("MLB200_TC_DEBUG")) : 0;
    p.halo = HaloMode(p.KC) ? 1 : 0;
    p.row_bytes = static_cast<int>(StageBytes(p.KC));
    int const want_rps = std::getenv("MLB200_TC_RPS") ? std::atoi(std::getenv("MLB200_TC_RPS")) : 3;
    // split output through staging + TMA store when four ring stages still fit next to it (thin layers: the
    // epilogue's direct 16-byte stores were the bound of the 3->64 layer, 3100 of its 3400 cycles per tile)
    p.split_stage = (out_split && d_out_bf16 != nullptr && EnvInt("MLB200_X3_STAGE", 1) != 0
                     && SmemBytes(p.N, Ct, p.KC, 4, 2 * p.N) <= 226u * 1024u) ? 1 : 0;
    int const staged_channels = out_bf16 ? p.N : (p.split_stage ? 2 * p.N : 0);
    p.rps = (want_rps == 3 && SmemBytes(p.N, Ct, p.KC, 2, staged_channels, 3) <= 226u * 1024u) ? 3 : 1;
    p.stage_bytes = p.rps * p.row_bytes;
    p.tap_stride = p.halo ? p.KC * 2 : kTileM * p.KC * 2;
    p.stage_tx = p.rps * (p.halo ? (kTileM + 2) * p.KC * 2 : 3 * kTileM * p.KC * 2);
    p.prof = nullptr;
    if (std::getenv("MLB200_TC_PROF") != nullptr)
        ML_CUDA_CHECK(cudaMalloc(reinterpret_cast<void**>(&p.prof), 16 * sizeof(unsigned long long)));
    p.rows_mode = (!split && !out_split && RowsModeFits(p.N, Cin, out_bf16)) ? 1 : 0;
    p.wait_ns = EnvInt("MLB200_TC_WAIT_NS", 0);
    p.stages = kMaxStages;
    if (p.rows_mode)
    {
        p.rps = 1;
        p.stage_bytes = p.row_bytes;
        p.stage_tx = p.halo ? (kTileM + 2) * p.KC * 2 : 3 * kTileM * p.KC * 2;
        p.n_acc = 2;                                  // two pipes of three slots
        cols = 32;
        while (cols < 6 * p.N) cols *= 2;
        p.tmem_cols = cols;
        while (p.stages > 3 && SmemBytesRows(p.N, p.KC, p.stages, out_bf16) > 226u * 1024u) p.stages--;
        op->smem = SmemBytesRows(p.N, p.KC, p.stages, out_bf16);
        int const want = (p.n_tiles + 1) / 2;        // two pipes per CTA
        op->grid = want < sm_count ? want : sm_count;
    }
    else
    {
        while (p.stages > 2 && SmemBytes(p.N, Ct, p.KC, p.stages, staged_channels, p.rps) > 226u * 1024u) p.stages--;
        op->smem = SmemBytes(p.N, Ct, p.KC, p.stages, staged_channels, p.rps);
        op->grid = p.n_tiles < sm_count ? p.n_tiles : sm_count;
    }
    if (op->smem > 226u * 1024u) throw std::runtime_error("filter does not fit in shared memory");

    // activation: [H][W][Cin] bf16, box = KC channels x 128 pixels x 1 row
    {
        cuuint64_t dims[3] = {static_cast<cuuint64_t>(Ct), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H)};
        cuuint64_t strides[2] = {static_cast<cuuint64_t>(Ct) * 2, static_cast<cuuint64_t>(W) * Ct * 2};
        cuuint32_t box[3] = {static_cast<cuuint32_t>(p.KC), static_cast<cuuint32_t>(p.halo ? kTileM + 2 : kTileM), 1};
        Encode(&op->in_map, d_in, 3, dims, strides, box, SwizzleFor(p.KC));
    }
    // filter: [9 * N][Cin] bf16, box = KC channels x N rows
    {
        // (wide split: [9 * 2N][Cin], box = KC channels x 2N rows -- the hi rows and the lo rows of a tap)
        int const Cw = p.wide ? Cin : Ct, rows = p.wide ? 2 * p.N : p.N;
        cuuint64_t dims[2] = {static_cast<cuuint64_t>(Cw), static_cast<cuuint64_t>(9) * rows};
        cuuint64_t strides[1] = {static_cast<cuuint64_t>(Cw) * 2};
        cuuint32_t box[2] = {static_cast<cuuint32_t>(p.KC), static_cast<cuuint32_t>(rows)};
        Encode(&op->w_map, d_w_packed, 2, dims, strides, box, SwizzleFor(p.KC));
    }
    // bf16 output: [H][W][N] bf16, box = min(N, 64) channels x 128 pixels (the TMA store clips at W)
    if (out_bf16 || p.split_stage)
    {
        if (d_out_bf16 == nullptr) throw std::runtime_error("bf16 output buffer missing");
        int const NO = p.split_stage ? 2 * p.N : p.N;      // split: [hi: N | lo: N] per pixel
        int const OC = NO < 64 ? NO : 64;
        cuuint64_t dims[3] = {static_cast<cuuint64_t>(NO), static_cast<cuuint64_t>(W), static_cast<cuuint64_t>(H)};
        cuuint64_t strides[2] = {static_cast<cuuint64_t>(NO) * 2, static_cast<cuuint64_t>(W) * NO * 2};
        cuuint32_t box[3] = {static_cast<cuuint32_t>(OC), kTileM, 1};
        Encode(&op->out_map, d_out_bf16, 3, dims, strides, box, SwizzleFor(OC));
    }
    else
    {
        op->out_map = op->in_map; // unused by the fp32-output epilogue
    }
    // per device: the attribute belongs to the function in the current context
    auto set = [](auto* k) { ML_CUDA_CHECK(cudaFuncSetAttribute(k, cudaFuncAttributeMaxDynamicSharedMemorySize, 226 * 1024)); };
    if (p.rows_mode) set(RowsKernelFor(p));
    else
    {
        if (p.KC == 64) set(conv3x3_tc_kernel<64>);
        else if (p.KC == 32) set(conv3x3_tc_kernel<32>);
        else set(conv3x3_tc_kernel<16>);
    }
    return op.release();
}

void ConvTcRun(ConvTc const* op, float const* bias, void* out, cudaStream_t stream)
{
    if (op->p.prof) ML_CUDA_CHECK(cudaMemsetAsync(op->p.prof, 0, 16 * sizeof(unsigned long long), stream));
    if (op->p.rows_mode)
    {
        RowsKernelFor(op->p)<<<op->grid, kTcThreads, op->smem, stream>>>(op->in_map, op->w_map, op->out_map, bias, out, op->p);
    }
    else if (op->p.KC == 64)
        conv3x3_tc_kernel<64><<<op->grid, kTcThreads, op->smem, stream>>>(op->in_map, op->w_map, op->out_map, bias, out, op->p);
    else if (op->p.KC == 32)
        conv3x3_tc_kernel<32><<<op->grid, kTcThreads, op->smem, stream>>>(op->in_map, op->w_map, op->out_map, bias, out, op->p);
    else
        conv3x3_tc_kernel<16><<<op->grid, kTcThreads, op->smem, stream>>>(op->in_map, op->w_map, op->out_map, bias, out, op->p);
    ML_CUDA_CHECK(cudaGetLastError());
    if (op->p.prof)
    {
        unsigned long long h[16];
        ML_CUDA_CHECK(cudaStreamSynchronize(stream));
        ML_CUDA_CHECK(cudaMemcpy(h, op->p.prof, sizeof(h), cudaMemcpyDeviceToHost));
        double const per = 1.0 / (static_cast<double>(op->grid) * ((op->p.n_tiles + op->grid - 1) / op->grid));
        std::fprintf(stderr, "[tc prof] N=%d rows_mode=%d cycles/tile: producer wait_empty %.0f other %.0f | mma wait_acc_empty %.0f "
                     "wait_full %.0f issue %.0f commit %.0f rest %.0f | epilogue wait_acc_full %.0f other %.0f ld+math+sts %.0f wait_store %.0f barrier %.0f\n", op->p.N, op->p.rows_mode,
                     h[0] * per, h[1] * per, h[2] * per, h[3] * per, h[4] * per, h[7] * per, h[8] * per, h[5] * per,
                     h[6] * per, h[9] * per, h[10] * per, h[11] * per);
        if (op->p.rows_mode)
            std::fprintf(stderr, "[tc prof]   epilogue detail: tcgen05.ld+wait %.0f (excluded from ld+math+sts above) | slot_free arrive %.0f fence.proxy.async %.0f\n",
                         h[12] * per, h[13] * per, h[14] * per);
    }
}

void ConvTcDestroy(ConvTc* op)
{
    if (op != nullptr) cudaFree(op->p.prof);
    delete op;
}

} // namespace nn
} // namespace ML
