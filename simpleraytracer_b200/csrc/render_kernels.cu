// render_kernels.cu -- the render loop as sm_100a CUDA kernels (SURVEY.md rows
// R1-R5; BUILD-DEFINED arithmetic, fixed by docs/RENDER_SPEC.md).
//
//  * render_brute_kernel<R>: every ray tests every triangle.  Triangles are
//    staged HBM -> shared memory as SoA float4 tiles with 1-D bulk async copies
//    (cp.async.bulk + mbarrier, double buffered); every thread keeps R rays in
//    registers and reads each triangle once per R tests as three broadcast
//    LDS.128.  This is the FP32-FMA-pipe roofline kernel (45 FLOP per test).
//  * render_bvh_kernel: one thread per pixel, stack traversal of the 64-byte
//    two-child-box nodes through the read-only path, near child first.
//  * render_bvh_spp_kernel<PIX>: a warp spreads the samples of PIX pixels over
//    its lanes (coherent rays), radiance still summed in sample order.
//  * grid_*_kernel + the grid paths inside the two BVH kernels: per-frame screen
//    grid (primary rays) and per-light perspective grid (shadow rays) that hand
//    every ray a short, exact candidate list instead of a tree walk
//    (docs/RENDER_SPEC.md section 8); the BVH stays as the fallback.
//
// All produce identical results: the winner of a closest-hit query is the
// lexicographic minimum of (t, id) and the barycentric decision is a pure
// function of (ray, triangle), so traversal order never shows in the output.
//
// Compiled with -fmad=false; every fused operation is an explicit __fmaf_rn.
#include "scene.h"

#include <cuda_runtime.h>

#include <algorithm>

#include <mutex>

namespace ML {
namespace {

// ---------------------------------------------------------------------------
// spec arithmetic (docs/RENDER_SPEC.md section 0)
// ---------------------------------------------------------------------------
struct V3 { float x, y, z; };

__device__ __forceinline__ float mul(float a, float b) { return __fmul_rn(a, b); }
__device__ __forceinline__ float add(float a, float b) { return __fadd_rn(a, b); }
__device__ __forceinline__ float sub(float a, float b) { return __fsub_rn(a, b); }
__device__ __forceinline__ float fma_(float a, float b, float c) { return __fmaf_rn(a, b, c); }

__device__ __forceinline__ float dot3(V3 a, V3 b)
{
    return fma_(a.z, b.z, fma_(a.y, b.y, mul(a.x, b.x)));
}
__device__ __forceinline__ V3 cross3(V3 a, V3 b)
{
    V3 r;
    r.x = fma_(a.y, b.z, -mul(a.z, b.y));
    r.y = fma_(a.z, b.x, -mul(a.x, b.z));
    r.z = fma_(a.x, b.y, -mul(a.y, b.x));
    return r;
}
__device__ __forceinline__ V3 v3(float const* p) { return V3{p[0], p[1], p[2]}; }
__device__ __forceinline__ V3 v3(float4 p) { return V3{p.x, p.y, p.z}; }
__device__ __forceinline__ float pos_inf() { return __int_as_float(0x7f800000); }

// Packed pair of fp32 in one 64-bit register (sm_100 f32x2 arithmetic).  Used
// only where results do not enter the image (BVH culling).
typedef unsigned long long F2;
__device__ __forceinline__ F2 pack2(float a, float b)
{
    F2 r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ F2 pack2(float a) { return pack2(a, a); }
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

__device__ __forceinline__ uint32_t hash_sample(uint32_t pix, uint32_t s, uint32_t dim, uint32_t seed)
{
    uint32_t h = (pix * 0x9E3779B1u) ^ (s * 0x85EBCA77u + dim * 0xC2B2AE3Du + seed * 0x27D4EB2Fu + 0x165667B1u);
    h ^= h >> 16; h *= 0x7FEB352Du; h ^= h >> 15; h *= 0x846CA68Bu; h ^= h >> 16;
    return h;
}
__device__ __forceinline__ float u01(uint32_t h) { return static_cast<float>(h >> 8) * (1.0f / 16777216.0f); }

// section 2: primary ray direction through pixel (x,y), sample s.
// `len_out`: |forward + sx*right + sy*up| of the sample, the factor between a point's depth coefficient along
// `forward` and its distance t along the unit direction (t = cz * len, exactly, for any invertible camera basis).
__device__ __forceinline__ V3 primary_dir(RenderLaunch const& L, uint32_t x, uint32_t y, uint32_t s, float* len_out = nullptr)
{
    float jx = 0.5f, jy = 0.5f;
    if (L.params.jitter)
    {
        uint32_t pix = y * L.params.width + x;
        jx = u01(hash_sample(pix, s, 0u, L.params.seed));
        jy = u01(hash_sample(pix, s, 1u, L.params.seed));
    }
    float fx = add(static_cast<float>(x), jx);
    float fy = add(static_cast<float>(y), jy);
    float sx = mul(fma_(fx, L.inv2w, -1.0f), L.camera.half_w);
    float sy = mul(-fma_(fy, L.inv2h, -1.0f), L.camera.half_h);
    V3 r;
    r.x = fma_(sy, L.camera.up[0], fma_(sx, L.camera.right[0], L.camera.forward[0]));
    r.y = fma_(sy, L.camera.up[1], fma_(sx, L.camera.right[1], L.camera.forward[1]));
    r.z = fma_(sy, L.camera.up[2], fma_(sx, L.camera.right[2], L.camera.forward[2]));
    float len = __fsqrt_rn(dot3(r, r));
    float inv = __fdiv_rn(1.0f, len);
    if (len_out != nullptr) *len_out = len;
    return V3{mul(r.x, inv), mul(r.y, inv), mul(r.z, inv)};
}

// section 3: sign-normalised Moller-Trumbore.  Returns the `inside` decision;
// ts/adet are the normalised numerator and |det| (t = ts/adet).
__device__ __forceinline__ bool mt_inside(V3 o, V3 d, float4 v0, float4 e1, float4 e2,
                                          float& ts, float& adet)
{
    V3 E1 = v3(e1), E2 = v3(e2);
    V3 p = cross3(d, E2);
    float det = dot3(E1, p);
    V3 tv = V3{sub(o.x, v0.x), sub(o.y, v0.y), sub(o.z, v0.z)};
    float us = dot3(tv, p);
    V3 q = cross3(tv, E1);
    float vs = dot3(d, q);
    float tsr = dot3(E2, q);
    uint32_t sgn = __float_as_uint(det) & 0x80000000u;
    us = __uint_as_float(__float_as_uint(us) ^ sgn);
    vs = __uint_as_float(__float_as_uint(vs) ^ sgn);
    ts = __uint_as_float(__float_as_uint(tsr) ^ sgn);
    adet = fabsf(det);
    // The spec's `det != 0` term is implied for every observable result and is
    // not evaluated: with det == 0, t = ts/0 is +-inf or NaN, which neither the
    // closest-hit rule (t < best.t, or t == best.t and id < best.id with
    // best.id = -1 while best.t = +inf) nor the any-hit rule (0 < t < 1) accepts.
    return (us >= 0.0f) & (vs >= 0.0f) & (add(us, vs) <= adet);
}

// Closest-hit record: t, original triangle id (tie-break + output) and the
// slot of the triangle in the leaf-ordered arrays (to fetch its edges later).
struct Hit { float t; int id; uint32_t slot; };

__device__ __forceinline__ Hit no_hit() { return Hit{pos_inf(), -1, 0u}; }

__device__ __forceinline__ void closest_update(Hit& best, float ts, float adet, int id, uint32_t slot,
                                               float t_min)
{
    float t = __fdiv_rn(ts, adet);
    if (t > t_min && (t < best.t || (t == best.t && id < best.id)))
    {
        best.t = t;
        best.id = id;
        best.slot = slot;
    }
}

// section 4 up to the shadow ray: the unshadowed radiance and, when a shadow
// ray is needed, its origin/direction.
struct ShadeOut { V3 rgb; V3 so, sd; bool need_shadow; };

// `n` is the triangle's geometric unit normal normalize(e1 x e2) -- a pure function of the triangle, evaluated
// once per triangle by tri_normals_kernel with the operation sequence of the spec (cross3, dot3, __fsqrt_rn,
// __fdiv_rn(1, len), three mul) instead of once per sample -- and `alb` its albedo.
__device__ __forceinline__ ShadeOut shade_hit(RenderLaunch const& L, V3 o, V3 d, float t, V3 n, V3 alb)
{
    ShadeOut out;
    out.need_shadow = false;
    out.so = V3{0.f, 0.f, 0.f};
    out.sd = V3{0.f, 0.f, 0.f};
    if (L.params.shade == SRT_SHADE_ALBEDO) { out.rgb = alb; return out; }

    V3 P = V3{fma_(t, d.x, o.x), fma_(t, d.y, o.y), fma_(t, d.z, o.z)};
    if (dot3(n, d) > 0.0f) n = V3{-n.x, -n.y, -n.z};
    V3 Lv = V3{sub(L.light.position[0], P.x), sub(L.light.position[1], P.y), sub(L.light.position[2], P.z)};
    float dist2 = dot3(Lv, Lv);
    float dist = __fsqrt_rn(dist2);
    float idist = __fdiv_rn(1.0f, dist);
    V3 l = V3{mul(Lv.x, idist), mul(Lv.y, idist), mul(Lv.z, idist)};
    float c = dot3(n, l);
    if (!(c > 0.0f)) { out.rgb = V3{0.f, 0.f, 0.f}; return out; }
    float k = __fdiv_rn(c, dist2);
    out.rgb = V3{mul(mul(alb.x, L.light.intensity[0]), k),
                 mul(mul(alb.y, L.light.intensity[1]), k),
                 mul(mul(alb.z, L.light.intensity[2]), k)};
    if (L.params.shade == SRT_SHADE_SHADOW)
    {
        out.need_shadow = true;
        float eps = L.params.shadow_eps;
        out.so = V3{fma_(eps, n.x, P.x), fma_(eps, n.y, P.y), fma_(eps, n.z, P.z)};
        out.sd = V3{sub(L.light.position[0], out.so.x), sub(L.light.position[1], out.so.y),
                    sub(L.light.position[2], out.so.z)};
    }
    return out;
}

// Shading inputs of a hit from the scene arrays in HBM / L2 (paths that do not hold them in shared memory).
__device__ __forceinline__ ShadeOut shade_hit(RenderLaunch const& L, V3 o, V3 d, Hit h)
{
    V3 const alb = v3(__ldg(reinterpret_cast<float4 const*>(L.scene.albedo) + h.id));
    V3 n = V3{0.f, 0.f, 0.f};
    if (L.params.shade != SRT_SHADE_ALBEDO) n = v3(__ldg(reinterpret_cast<float4 const*>(L.scene.tri_n) + h.slot));
    return shade_hit(L, o, d, h.t, n, alb);
}

__device__ __forceinline__ unsigned long long warp_sum(unsigned long long v)
{
    for (int off = 16; off > 0; off >>= 1) v += __shfl_xor_sync(0xffffffffu, v, off);
    return v;
}

// End of a render block: ray counters and the completion signal.
//  * Counters: warp sums -> shared-memory atomics -> ONE global atomic per counter per block (and only for
//    counters that are non-zero); box / triangle test counts exist only in the collect_stats variants.
//  * Signal (RenderLaunch::d_signal, multi-GPU frames): thread 0 of every block fences its block's stores at
//    system scope (they may be peer stores into another GPU's frame) and takes a ticket; the block that draws
//    the last ticket resets it and publishes signal_value with a system-scope release store.  A consumer that
//    observes the value with an acquire load (wait_flags_kernel) therefore sees every pixel of the launch.
// Called by every thread of the block, outside divergent control flow.
__device__ __forceinline__ void flush_counters(RenderLaunch const& L, unsigned long long primary,
                                               unsigned long long shadow, unsigned long long boxes,
                                               unsigned long long tris)
{
    __shared__ unsigned long long s_cnt[4];
    if (threadIdx.x < 4) s_cnt[threadIdx.x] = 0ull;
    __syncthreads();
    primary = warp_sum(primary);
    shadow = warp_sum(shadow);
    if (L.params.collect_stats)
    {
        boxes = warp_sum(boxes);
        tris = warp_sum(tris);
    }
    if ((threadIdx.x & 31) == 0)
    {
        if (primary) atomicAdd(&s_cnt[0], primary);
        if (shadow) atomicAdd(&s_cnt[1], shadow);
        if (L.params.collect_stats)
        {
            atomicAdd(&s_cnt[2], boxes);
            atomicAdd(&s_cnt[3], tris);
        }
    }
    __syncthreads();   // also orders every thread's pixel stores before thread 0's fence below
    if (threadIdx.x == 0)
    {
#pragma unroll
        for (int k = 0; k < 4; k++)
            if (s_cnt[k]) atomicAdd(L.d_counters + k, s_cnt[k]);
        if (L.d_signal)
        {
            __threadfence_system();
            uint32_t const blocks = gridDim.x * gridDim.y * gridDim.z;
            if (atomicAdd(L.d_ticket, 1u) == blocks - 1u)
            {
                *L.d_ticket = 0u;   // next launch on this stream starts from zero
                __threadfence_system();
                asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(L.d_signal), "r"(L.signal_value) : "memory");
            }
        }
    }
}

// Block timeline probe (SRT_BLOCK_TRACE=1): see RenderLaunch::d_trace.
__device__ __forceinline__ unsigned long long trace_now()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}
__device__ __forceinline__ void trace_block(RenderLaunch const& L, unsigned long long t_start, uint32_t tile = 0,
                                            uint32_t cnt = 0, unsigned long long d_box = 0, unsigned long long d_tri = 0,
                                            uint32_t d_shadow = 0)
{
    if (L.d_trace == nullptr || threadIdx.x != 0) return;
    unsigned long long const t_end = trace_now();
    // longest item with its identity: duration ns << 34 | min(list length, 1023) << 24 | tile index; and (racy,
    // diagnostic only) what thread 0 did in it: box tests, triangle tests, shadow rays (collect_stats variant)
    unsigned long long const key = ((t_end - t_start) << 34) | (static_cast<unsigned long long>(min(cnt, 1023u)) << 24)
                                   | (tile & 0xffffffu);
    if (key > atomicMax(L.d_trace + 5, key))
    {
        L.d_trace[6] = d_box;
        L.d_trace[7] = d_tri | (static_cast<unsigned long long>(d_shadow) << 40);
    }
    atomicMax(L.d_trace + 0, ~t_start);          // ~min(start)
    atomicMax(L.d_trace + 1, t_start);
    atomicMax(L.d_trace + 2, t_end);
    atomicAdd(L.d_trace + 3, t_end - t_start);
    atomicMax(L.d_trace + 4, t_end - t_start);
}

// Writes one finished pixel (section 5).  `pix` is the full-frame pixel index.
__device__ __forceinline__ void store_pixel(RenderLaunch const& L, uint32_t pix, V3 acc, Hit first)
{
    size_t const out = static_cast<size_t>(pix) - static_cast<size_t>(L.params.buf_y0) * L.params.width;
    if (L.d_rgb)
    {
        float* p = L.d_rgb + out * 3;
        p[0] = mul(acc.x, L.inv_spp);
        p[1] = mul(acc.y, L.inv_spp);
        p[2] = mul(acc.z, L.inv_spp);
    }
    if (L.d_hit_id) L.d_hit_id[out] = first.id;
    if (L.d_hit_t) L.d_hit_t[out] = first.t;
}

// ---------------------------------------------------------------------------
// mbarrier / bulk-copy helpers (PTX)
// ---------------------------------------------------------------------------
__device__ __forceinline__ uint32_t smem_u32(void const* p)
{
    return static_cast<uint32_t>(__cvta_generic_to_shared(p));
}
__device__ __forceinline__ void mbar_init(uint64_t* bar, uint32_t count)
{
    asm volatile("mbarrier.init.shared::cta.b64 [%0], %1;" ::"r"(smem_u32(bar)), "r"(count) : "memory");
}
__device__ __forceinline__ void mbar_expect_tx(uint64_t* bar, uint32_t bytes)
{
    asm volatile("mbarrier.arrive.expect_tx.shared::cta.b64 _, [%0], %1;" ::"r"(smem_u32(bar)), "r"(bytes)
                 : "memory");
}
__device__ __forceinline__ void mbar_wait(uint64_t* bar, uint32_t parity)
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
            : "r"(smem_u32(bar)), "r"(parity)
            : "memory");
    } while (!done);
}
__device__ __forceinline__ void bulk_g2s(void* dst_smem, void const* src_gmem, uint32_t bytes, uint64_t* bar)
{
    asm volatile(
        "cp.async.bulk.shared::cluster.global.mbarrier::complete_tx::bytes [%0], [%1], %2, [%3];"
        ::"r"(smem_u32(dst_smem)), "l"(src_gmem), "r"(bytes), "r"(smem_u32(bar))
        : "memory");
}

// ---------------------------------------------------------------------------
// brute force
// ---------------------------------------------------------------------------
constexpr int kBruteThreads = 256;
constexpr uint32_t kTileBytes = kTriTile * sizeof(float4);

struct BruteSmem
{
    float4 v0[2][kTriTile];
    float4 e1[2][kTriTile];
    float4 e2[2][kTriTile];
    uint64_t bar[2];
};

__device__ __forceinline__ void brute_issue_tile(BruteSmem& sm, DeviceSceneView const& sc,
                                                 uint32_t tile, uint32_t buf)
{
    mbar_expect_tx(&sm.bar[buf], 3u * kTileBytes);
    size_t off = static_cast<size_t>(tile) * kTriTile;
    bulk_g2s(sm.v0[buf], sc.tri_v0 + off, kTileBytes, &sm.bar[buf]);
    bulk_g2s(sm.e1[buf], sc.tri_e1 + off, kTileBytes, &sm.bar[buf]);
    bulk_g2s(sm.e2[buf], sc.tri_e2 + off, kTileBytes, &sm.bar[buf]);
}

// One block-collective sweep over all triangle tiles.  ANY = false: closest
// hit for R rays; ANY = true: occlusion flags.  `seq` counts tiles consumed by
// this block so far (buffer = seq & 1, mbarrier parity = (seq >> 1) & 1).
template <int R, bool ANY>
__device__ __forceinline__ void brute_sweep(BruteSmem& sm, DeviceSceneView const& sc, uint32_t& seq,
                                            V3 const (&o)[R], V3 const (&d)[R], float t_min,
                                            Hit (&best)[R], bool (&occluded)[R])
{
    uint32_t const n_tiles = sc.n_tris_padded / kTriTile;
    if (threadIdx.x == 0)
    {
        brute_issue_tile(sm, sc, 0, seq & 1u);
        if (n_tiles > 1) brute_issue_tile(sm, sc, 1, (seq + 1u) & 1u);
    }
    for (uint32_t tile = 0; tile < n_tiles; tile++, seq++)
    {
        uint32_t const buf = seq & 1u;
        mbar_wait(&sm.bar[buf], (seq >> 1) & 1u);
        float4 const* tv0 = sm.v0[buf];
        float4 const* te1 = sm.e1[buf];
        float4 const* te2 = sm.e2[buf];
#pragma unroll 2
        for (int j = 0; j < kTriTile; j++)
        {
            float4 v0 = tv0[j], e1 = te1[j], e2 = te2[j];
            // All R barycentric decisions first (straight-line, R independent
            // dependency chains), then ONE rarely-taken branch for the division
            // and the (t, id) update: hits are a ~1e-5 fraction of the tests.
            float ts[R], adet[R];
            bool in[R];
            bool any = false;
#pragma unroll
            for (int r = 0; r < R; r++)
            {
                in[r] = mt_inside(o[r], d[r], v0, e1, e2, ts[r], adet[r]);
                any |= in[r];
            }
            if (any)
            {
#pragma unroll
                for (int r = 0; r < R; r++)
                {
                    if (!in[r]) continue;
                    if (ANY)
                    {
                        float t = __fdiv_rn(ts[r], adet[r]);
                        if (t > 0.0f && t < 1.0f) occluded[r] = true;
                    }
                    else
                    {
                        closest_update(best[r], ts[r], adet[r], __float_as_int(v0.w),
                                       tile * kTriTile + static_cast<uint32_t>(j), t_min);
                    }
                }
            }
        }
        __syncthreads(); // everyone is done with `buf` before it is refilled
        if (threadIdx.x == 0 && tile + 2 < n_tiles) brute_issue_tile(sm, sc, tile + 2, buf);
    }
}

template <int R>
__global__ void __launch_bounds__(kBruteThreads) render_brute_kernel(const __grid_constant__ RenderLaunch L)
{
    __shared__ __align__(128) BruteSmem sm;
    if (threadIdx.x == 0)
    {
        mbar_init(&sm.bar[0], 1);
        mbar_init(&sm.bar[1], 1);
        asm volatile("fence.mbarrier_init.release.cluster;" ::: "memory");
    }
    __syncthreads();

    uint32_t const W = L.params.width;
    uint32_t const owned_pixels = L.n_row_blocks * 8u * W; // rows past y1 are clipped below
    uint32_t const chunk = blockIdx.x * (kBruteThreads * R);
    V3 const cam_o = v3(L.camera.origin);
    DeviceSceneView const& sc = L.scene;
    bool const shadows = L.params.shade == SRT_SHADE_SHADOW;

    uint32_t pix[R];
    bool valid[R];
    V3 acc[R];
    Hit first_hit[R];
#pragma unroll
    for (int r = 0; r < R; r++)
    {
        uint32_t local = chunk + r * kBruteThreads + threadIdx.x;
        // local pixel -> (owned row, x) -> frame row through the row-block map
        uint32_t lrow = local / W, lx = local - lrow * W;
        uint32_t fy = L.row_origin + (lrow >> 3) * L.row_block_step + (lrow & 7u);
        valid[r] = local < owned_pixels && fy < L.params.y1;
        pix[r] = valid[r] ? fy * W + lx : L.params.y0 * W;
        acc[r] = V3{0.f, 0.f, 0.f};
        first_hit[r] = no_hit();
    }

    uint32_t seq = 0;
    unsigned long long n_primary = 0, n_shadow = 0;

    for (uint32_t s = 0; s < L.params.spp; s++)
    {
        V3 o[R], d[R];
        Hit best[R];
        bool occ[R];
#pragma unroll
        for (int r = 0; r < R; r++)
        {
            uint32_t y = pix[r] / W, x = pix[r] - y * W;
            o[r] = cam_o;
            // Lanes past the end of the band trace a null ray (d = 0 => det = 0
            // => never inside) so the block-collective sweep stays uniform.
            d[r] = valid[r] ? primary_dir(L, x, y, s) : V3{0.f, 0.f, 0.f};
            best[r] = no_hit();
            occ[r] = false;
            n_primary += valid[r] ? 1u : 0u;
        }
        brute_sweep<R, false>(sm, sc, seq, o, d, L.params.t_min, best, occ);

        V3 rgb[R];
#pragma unroll
        for (int r = 0; r < R; r++)
        {
            if (s == 0) first_hit[r] = best[r];
            V3 so = V3{0.f, 0.f, 0.f}, sd = V3{0.f, 0.f, 0.f};
            if (best[r].id < 0 || !valid[r])
            {
                rgb[r] = v3(L.params.background);
            }
            else
            {
                ShadeOut sh = shade_hit(L, o[r], d[r], best[r]);
                rgb[r] = sh.rgb;
                if (sh.need_shadow)
                {
                    so = sh.so;
                    sd = sh.sd;
                    n_shadow += 1u;
                }
            }
            o[r] = so;
            d[r] = sd; // null ray when no shadow query is needed
        }
        if (shadows) // block-uniform
        {
            brute_sweep<R, true>(sm, sc, seq, o, d, 0.0f, best, occ);
#pragma unroll
            for (int r = 0; r < R; r++)
                if (occ[r]) rgb[r] = V3{0.f, 0.f, 0.f};
        }
#pragma unroll
        for (int r = 0; r < R; r++)
        {
            acc[r].x = add(acc[r].x, rgb[r].x);
            acc[r].y = add(acc[r].y, rgb[r].y);
            acc[r].z = add(acc[r].z, rgb[r].z);
        }
    }

#pragma unroll
    for (int r = 0; r < R; r++)
        if (valid[r]) store_pixel(L, pix[r], acc[r], first_hit[r]);

    unsigned long long tri_tests = (n_primary + n_shadow) * sc.n_tris;
    flush_counters(L, n_primary, n_shadow, 0ull, tri_tests);
}

// ---------------------------------------------------------------------------
// Primary-ray screen grid (scene.h GridView): count / scan / fill, rebuilt per render call
// ---------------------------------------------------------------------------
// Camera-space coefficients of P: P - origin = cx*right + cy*up + cz*forward.  A primary ray through
// the sample (sx, sy) has direction forward + sx*right + sy*up, so it passes through P exactly when
// cx/cz = sx and cy/cz = sy (cz > 0).  Every point of a triangle whose three vertices have cz > 0
// projects inside the bounding box of the projected vertices.  The box is dilated by the projection
// of a world-space slack (4x the BVH box padding, plus the fp32 error of the projection itself) and
// by 1/4 pixel, which exceeds by orders of magnitude how far outside a triangle the fp32
// Moller-Trumbore test can still answer "inside" (the BVH boxes rely on a quarter of that slack).
// edge[k] = (nx, ny, c, *): nx*q.x + ny*q.y + c >= 0 for every query q (centre of a pixel square, or a point)
// whose rays can hit the triangle; (0, 0, 1, *) = no statement.  edge[0].w = nearest vertex depth minus slack.
__device__ __forceinline__ bool grid_project(GridBuild const& G, uint32_t i, float4& box, float4 (&edge)[3])
{
    edge[0] = edge[1] = edge[2] = make_float4(0.f, 0.f, 1.f, 0.f);
    float4 const v0 = reinterpret_cast<float4 const*>(G.scene.tri_v0)[i];
    float4 const e1 = reinterpret_cast<float4 const*>(G.scene.tri_e1)[i];
    float4 const e2 = reinterpret_cast<float4 const*>(G.scene.tri_e2)[i];
    float tx0 = pos_inf(), tx1 = -pos_inf(), ty0 = pos_inf(), ty1 = -pos_inf();
    float zmin = pos_inf(), zmax = -pos_inf(), dist = 0.f, tmax = 0.f;
    float vx[3] = {v0.x, v0.x + e1.x, v0.x + e2.x};
    float vy[3] = {v0.y, v0.y + e1.y, v0.y + e2.y};
    float vz[3] = {v0.z, v0.z + e1.z, v0.z + e2.z};
    float cxs[3], cys[3], czs[3];
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        float const px = vx[k] - G.origin[0], py = vy[k] - G.origin[1], pz = vz[k] - G.origin[2];
        cxs[k] = G.minv[0] * px + G.minv[1] * py + G.minv[2] * pz;
        cys[k] = G.minv[3] * px + G.minv[4] * py + G.minv[5] * pz;
        czs[k] = G.minv[6] * px + G.minv[7] * py + G.minv[8] * pz;
        zmin = fminf(zmin, czs[k]);
        zmax = fmaxf(zmax, czs[k]);
        dist = fmaxf(dist, fabsf(px) + fabsf(py) + fabsf(pz));
    }
    // basis scale: |c| <= mnorm * |P - o|
    float mnorm = 0.f;
#pragma unroll
    for (int k = 0; k < 9; k++) mnorm = fmaxf(mnorm, fabsf(G.minv[k]));
    float const slack_w = (G.pad + 4e-6f * dist) * 3.0f * mnorm;   // camera-space slack of each coefficient
    box = make_float4(1.f, 1.f, 0.f, 0.f);                         // x0 > x1: not binned
    if (zmax < -slack_w) return false;                             // wholly behind the camera: no ray reaches it
    if (!(zmin > 1e-3f * mnorm * dist + 8.0f * slack_w)) return true;   // near or crossing the camera plane: global
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        float const iz = 1.0f / czs[k];
        float const tx = cxs[k] * iz, ty = cys[k] * iz;
        tx0 = fminf(tx0, tx); tx1 = fmaxf(tx1, tx);
        ty0 = fminf(ty0, ty); ty1 = fmaxf(ty1, ty);
        tmax = fmaxf(tmax, fmaxf(fabsf(tx), fabsf(ty)));
    }
    // d(c/z) <= dc/z + |c| dz / z^2, doubled: the query point of the light grid is projected in fp32 too
    float const slack_t = 4.0f * slack_w * (1.0f + tmax) / zmin;
    float const kx = 0.5f * static_cast<float>(G.width) / G.half_w, ky = 0.5f * static_cast<float>(G.height) / G.half_h;
    float const cxp = 0.5f * static_cast<float>(G.width), cyp = 0.5f * static_cast<float>(G.height);
    // pixel x = (sx / half_w + 1) * W/2 ; pixel y = (1 - sy / half_h) * H/2
    float const x0 = cxp + (tx0 - slack_t) * kx - 0.25f, x1 = cxp + (tx1 + slack_t) * kx + 0.25f;
    float const y0 = cyp - (ty1 + slack_t) * ky - 0.25f, y1 = cyp - (ty0 - slack_t) * ky + 0.25f;
    if (!(x0 <= x1) || !(y0 <= y1) || !isfinite(x0) || !isfinite(x1) || !isfinite(y0) || !isfinite(y1)) return true;
    box = make_float4(x0, y0, x1, y1);
    edge[0].w = zmin - slack_w;
    // Edge functions of the projected triangle (perspective keeps a triangle in front of the plane a
    // triangle).  Margin: the dilation of the box (slack + 1/4 unit) plus the half extent of a query,
    // in the L1 norm of the edge normal -- the largest value the function takes over the dilated
    // square around a query centre.  Skipped for slivers and for coordinates whose products would
    // round coarser than the margin.
    float qx[3], qy[3], big = 0.f;
#pragma unroll
    for (int k = 0; k < 3; k++)
    {
        float const iz = 1.0f / czs[k];
        qx[k] = cxp + cxs[k] * iz * kx;
        qy[k] = cyp - cys[k] * iz * ky;
        big = fmaxf(big, fmaxf(fabsf(qx[k]), fabsf(qy[k])));
    }
    float const area2 = (qx[1] - qx[0]) * (qy[2] - qy[0]) - (qy[1] - qy[0]) * (qx[2] - qx[0]);
    float const ext = fmaxf(x1 - x0, y1 - y0);
    // fp32 budget: coordinates carry an absolute error of ~1e-7 * big, so an edge normal is off by that
    // much and the function by up to that times the distance to the query (<= ext); the orientation
    // (sign of area2) must be beyond doubt, and each edge is used only if a quarter of a unit of its
    // margin covers these errors several times over.
    if (big < 2e5f && fabsf(area2) > 1e-5f * big * ext)
    {
        float const sgn = area2 > 0.f ? 1.f : -1.f;
        float const m = slack_t * fmaxf(kx, ky) + 0.25f + G.half_extent;
#pragma unroll
        for (int k = 0; k < 3; k++)
        {
            int const a = k, b = (k + 1) % 3;
            float const nx = -sgn * (qy[b] - qy[a]), ny = sgn * (qx[b] - qx[a]);
            float const n1 = fabsf(nx) + fabsf(ny);
            if (!(n1 >= 3.2e-6f * big * ext)) continue;
            float const c = -(nx * qx[a] + ny * qy[a]) + m * n1;
            edge[k].x = nx; edge[k].y = ny; edge[k].z = c;
        }
    }
    return false;
}

// tiles a box touches; false when it lies off the frame
__device__ __forceinline__ bool grid_tile_range(GridBuild const& G, float4 box, int& tx0, int& tx1, int& ty0, int& ty1)
{
    float const inv = 1.0f / static_cast<float>(G.tile);
    float const fx0 = floorf((box.x - G.lo_bias) * inv), fx1 = floorf(box.z * inv);
    float const fy0 = floorf((box.y - G.lo_bias) * inv), fy1 = floorf(box.w * inv);
    if (fx1 < 0.f || fy1 < 0.f || fx0 >= static_cast<float>(G.tiles_w) || fy0 >= static_cast<float>(G.tiles_h)) return false;
    tx0 = static_cast<int>(fmaxf(fx0, 0.f)); tx1 = static_cast<int>(fminf(fx1, static_cast<float>(G.tiles_w - 1)));
    ty0 = static_cast<int>(fmaxf(fy0, 0.f)); ty1 = static_cast<int>(fminf(fy1, static_cast<float>(G.tiles_h - 1)));
    return true;
}


// Does this build own tile row ty (GridBuild::ty_*)?
__device__ __forceinline__ bool grid_row_owned(GridBuild const& G, int ty)
{
    uint32_t const t = static_cast<uint32_t>(ty);
    return t >= G.ty_begin && t < G.ty_end && (G.ty_stride <= 1u || t % G.ty_stride == G.ty_phase);
}

// Can any query inside tile (tx, ty) pass the triangle's edge functions?  The functions already carry the
// margin of ONE query (half_extent); a tile is a square of `tile` units, so its centre gets the remaining
// half width in the L1 norm of each normal.  Evaluated identically by the count and the fill kernel.
__device__ __forceinline__ bool grid_tile_touches(GridBuild const& G, float4 const (&edge)[3], int tx, int ty)
{
    float const t = static_cast<float>(G.tile);
    float const qx = (static_cast<float>(tx) + 0.5f) * t, qy = (static_cast<float>(ty) + 0.5f) * t;
    float const h = 0.5f * t - G.half_extent;
    bool ok = true;
#pragma unroll
    for (int k = 0; k < 3; k++)
        ok = ok && (edge[k].x * qx + edge[k].y * qy + edge[k].z + h * (fabsf(edge[k].x) + fabsf(edge[k].y)) >= 0.0f);
    return ok;
}

// The tile rows of [ty0, ty1] this build owns (GridBuild::ty_*): `rows` of them, the first one `fy`, `step` apart.
// The walk below visits ONLY these: at 8 ranks a triangle's box holds an owned row about once in eight, and testing
// every tile of every box for ownership (a modulo per tile) was most of the count and fill kernels' instructions.
__device__ __forceinline__ void grid_owned_rows(GridBuild const& G, int ty0, int ty1, int& fy, uint32_t& rows)
{
    int lo = max(ty0, static_cast<int>(G.ty_begin));
    int const hi = min(ty1, static_cast<int>(G.ty_end) - 1);
    if (G.ty_stride > 1u)
    {
        if (G.ty_phase >= G.ty_stride) { fy = 0; rows = 0; return; }
        uint32_t const m = static_cast<uint32_t>(lo) % G.ty_stride;
        lo += static_cast<int>((G.ty_phase + G.ty_stride - m) % G.ty_stride);
    }
    fy = lo;
    rows = hi >= lo ? static_cast<uint32_t>(hi - lo) / max(G.ty_stride, 1u) + 1u : 0u;
}

// Owned tiles of one triangle's box, visited by the calling thread alone (COOP = false: row by row, no division)
// or by the 32 lanes of its warp together (COOP = true: lane t takes tiles t, t + 32, ...).  A thread per
// triangle serialises the few triangles that cover hundreds of tiles -- up to max_tiles_per_tri dependent atomics
// in one thread, which was most of the build time of frames with large triangles; the warp walks those boxes 32
// tiles at a time.  The threshold is not lower because the warp walks its large boxes ONE AFTER ANOTHER (~100
// instructions of broadcast and loop each, at one warp's issue rate): with 16, the slowest warp of the benchmark
// frame (24 boxes of 17..42 tiles) set the duration of the whole launch.
constexpr uint32_t kGridCoopTiles = 32;   // boxes with more owned tiles are walked by the whole warp

template <bool COOP, class Visit>
__device__ __forceinline__ void grid_visit_tiles(GridBuild const& G, float4 const (&edge)[3], int tx0, int tx1, int fy,
                                                 uint32_t rows, Visit&& visit)
{
    int const step = static_cast<int>(max(G.ty_stride, 1u));
    if (COOP)
    {
        uint32_t const w = static_cast<uint32_t>(tx1 - tx0 + 1), n = w * rows;
        for (uint32_t t = threadIdx.x & 31u; t < n; t += 32u)
        {
            uint32_t const r = t / w;
            int const ty = fy + static_cast<int>(r) * step, tx = tx0 + static_cast<int>(t - r * w);
            if (grid_tile_touches(G, edge, tx, ty)) visit(static_cast<uint32_t>(ty) * G.tiles_w + tx);
        }
    }
    else
    {
        int ty = fy;
        for (uint32_t r = 0; r < rows; r++, ty += step)
            for (int tx = tx0; tx <= tx1; tx++)
                if (grid_tile_touches(G, edge, tx, ty)) visit(static_cast<uint32_t>(ty) * G.tiles_w + tx);
    }
}

// Walks the boxes of a warp's triangles: small ones per thread, then every large one by the whole warp (its box,
// edge functions and slot broadcast from the owning lane).  Called by all 32 lanes.
template <class Visit>
__device__ __forceinline__ void grid_walk_warp(GridBuild const& G, bool binned, uint32_t slot, float4 box,
                                               float4 const (&edge)[3], int tx0, int tx1, int ty0, int ty1, Visit&& visit)
{
    int fy = 0;
    uint32_t rows = 0;
    if (binned) grid_owned_rows(G, ty0, ty1, fy, rows);
    uint32_t const n = binned ? static_cast<uint32_t>(tx1 - tx0 + 1) * rows : 0u;
    bool const big = n > kGridCoopTiles;
    if (binned && !big)
        grid_visit_tiles<false>(G, edge, tx0, tx1, fy, rows, [&](uint32_t tile) { visit(tile, slot, box); });
    uint32_t todo = __ballot_sync(0xffffffffu, big);
    while (todo)
    {
        int const src = __ffs(todo) - 1;
        todo &= todo - 1u;
        float4 e[3];
#pragma unroll
        for (int k = 0; k < 3; k++)
        {
            e[k].x = __shfl_sync(0xffffffffu, edge[k].x, src);
            e[k].y = __shfl_sync(0xffffffffu, edge[k].y, src);
            e[k].z = __shfl_sync(0xffffffffu, edge[k].z, src);
            e[k].w = 0.f;
        }
        int const bx0 = __shfl_sync(0xffffffffu, tx0, src), bx1 = __shfl_sync(0xffffffffu, tx1, src);
        int const bfy = __shfl_sync(0xffffffffu, fy, src);
        uint32_t const brows = __shfl_sync(0xffffffffu, rows, src);
        uint32_t const bslot = __shfl_sync(0xffffffffu, slot, src);
        float4 const bbox = make_float4(__shfl_sync(0xffffffffu, box.x, src), __shfl_sync(0xffffffffu, box.y, src),
                                        __shfl_sync(0xffffffffu, box.z, src), __shfl_sync(0xffffffffu, box.w, src));
        grid_visit_tiles<true>(G, e, bx0, bx1, bfy, brows, [&](uint32_t tile) { visit(tile, bslot, bbox); });
    }
}

// Phase bodies of the grid build (`vb` = block index; called by all threads of the block).  Negative result kept for
// the record: all phases in ONE cooperative launch with counter barriers between them was not faster than one launch
// per phase (config 3: 95 vs 91 us, config 4: 318 vs 243 us, round-2 figures before the walk below was rewritten).
// What the build's time was made of, measured per launch (ncu, warm, one rank's share of the 100k-triangle frame at
// 8 ranks): count 20 us + fill 30 us, both set by ONE warp -- the one whose triangles had 17..42-tile boxes, walked one
// after the other by the whole warp, with an ownership modulo and a division per tile.  Visiting owned rows only, row
// by row, brought them to 6.9 + 6.8 us (whole build with the scan changes below: 71 -> 31 us at 8 ranks, 75 -> 56 us on one GPU).
__device__ __forceinline__ void grid_phase_count(GridBuild const& G, uint32_t vb)
{
    uint32_t const i = vb * blockDim.x + threadIdx.x;     // n_tris_padded is a multiple of the block size
    float4 box = make_float4(1.f, 1.f, 0.f, 0.f);
    float4 edge[3] = {make_float4(0.f, 0.f, 1.f, 0.f), make_float4(0.f, 0.f, 1.f, 0.f), make_float4(0.f, 0.f, 1.f, 0.f)};
    bool global = false;
    bool const real = i < G.scene.n_tris && __float_as_int(reinterpret_cast<float4 const*>(G.scene.tri_v0)[i].w) >= 0;
    if (real) global = grid_project(G, i, box, edge);
    int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
    bool binned = !global && box.x <= box.z && grid_tile_range(G, box, tx0, tx1, ty0, ty1);
    if (binned && static_cast<uint32_t>((tx1 - tx0 + 1) * (ty1 - ty0 + 1)) > G.max_tiles_per_tri) global = true;
    if (global)
    {
        uint32_t const k = atomicAdd(&G.flags->n_global, 1u);
        if (k < kGridMaxGlobal) G.global[k] = i;
        binned = false;
    }
    if (!binned) box = make_float4(1.f, 1.f, 0.f, 0.f);   // global, off the frame or behind the camera
    grid_walk_warp(G, binned, i, box, edge, tx0, tx1, ty0, ty1,
                   [&](uint32_t tile, uint32_t, float4) { atomicAdd(&G.tile_cnt[tile], 1u); });
    if (i < G.scene.n_tris_padded)
    {
        reinterpret_cast<float4*>(G.bbox)[i] = box;
#pragma unroll
        for (int k = 0; k < 3; k++) reinterpret_cast<float4*>(G.edges)[3 * i + k] = edge[k];
    }
}

// Exclusive prefix sum of the tile counts in two small launches (the light grid has ~10^6 cells): per-block sums of
// 1024 cells -- the block that finishes last scans those sums and writes the verdict -- then per-block scans.  Small
// blocks on purpose: the frame's 32 400 tiles spread over 32 SMs instead of 8 (a per-block scan is a chain of
// dependent global loads, a shared-memory scan and one atomic per class: latency, which more SMs hide).
constexpr uint32_t kScanBlock = 256, kScanPerThread = 4, kScanCells = kScanBlock * kScanPerThread;
static_assert(kScanCells == kGridScanCells, "GridBuild::block_sums is sized by kGridScanCells");

__device__ __forceinline__ uint32_t block_exclusive_scan(uint32_t v, uint32_t* s_warp, uint32_t& block_total)
{
    uint32_t const lane = threadIdx.x & 31u, warp = threadIdx.x >> 5;
    uint32_t inc = v;
#pragma unroll
    for (uint32_t off = 1; off < 32u; off <<= 1)
    {
        uint32_t const n = __shfl_up_sync(0xffffffffu, inc, off);
        if (lane >= off) inc += n;
    }
    if (lane == 31u) s_warp[warp] = inc;
    __syncthreads();
    if (warp == 0)
    {
        uint32_t w = lane < (blockDim.x >> 5) ? s_warp[lane] : 0u;
        uint32_t winc = w;
#pragma unroll
        for (uint32_t off = 1; off < 32u; off <<= 1)
        {
            uint32_t const n = __shfl_up_sync(0xffffffffu, winc, off);
            if (lane >= off) winc += n;
        }
        s_warp[lane] = winc - w;           // exclusive offset of each warp
        if (lane == 31u) s_warp[32] = winc; // block total
    }
    __syncthreads();
    block_total = s_warp[32];
    return s_warp[warp] + inc - v;
}

__device__ __forceinline__ void grid_phase_scan_sums(GridBuild const& G, uint32_t n_blocks, uint32_t* s_warp)
{
    uint32_t const per = (n_blocks + kScanBlock - 1u) / kScanBlock;
    uint32_t const b = threadIdx.x * per, e = min(n_blocks, b + per);
    uint32_t sum = 0;
    for (uint32_t t = b; t < e; t++) sum += __ldcg(G.block_sums + t);
    uint32_t total;
    uint32_t run = block_exclusive_scan(sum, s_warp, total);
    for (uint32_t t = b; t < e; t++)
    {
        uint32_t const v = __ldcg(G.block_sums + t);
        G.block_sums[t] = run;
        run += v;
    }
    if (threadIdx.x == 0)
    {
        G.tile_off[G.tiles_w * G.tiles_h] = total;
        G.flags->total = total;
        G.flags->ok = (total <= G.cap && G.flags->n_global <= kGridMaxGlobal) ? 1u : 0u;
    }
}

__device__ __forceinline__ void grid_phase_block_sums(GridBuild const& G, uint32_t vb, uint32_t n_blocks)
{
    __shared__ uint32_t s_warp[33];
    uint32_t const tiles = G.tiles_w * G.tiles_h;
    uint32_t const base = vb * kScanCells + threadIdx.x * kScanPerThread;
    uint32_t sum = 0;
#pragma unroll
    for (uint32_t k = 0; k < kScanPerThread; k++)
        if (base + k < tiles) sum += G.tile_cnt[base + k];
    uint32_t total;
    block_exclusive_scan(sum, s_warp, total);
    // The block that finishes last scans the block sums (and writes the verdict): one launch less than a kernel of
    // its own.  Ticket pattern: publish, fence, count; the last arrival fences again and reads through L2.
    __shared__ bool s_last;
    if (threadIdx.x == 0)
    {
        G.block_sums[vb] = total;
        __threadfence();
        s_last = atomicAdd(&G.flags->ticket, 1u) == n_blocks - 1u;
    }
    __syncthreads();
    if (!s_last) return;
    __threadfence();
    grid_phase_scan_sums(G, n_blocks, s_warp);
}

__device__ __forceinline__ void grid_phase_scan_blocks(GridBuild const& G, uint32_t vb)
{
    __shared__ uint32_t s_warp[33];
    __shared__ uint32_t s_cls[kGridBuckets], s_cls_base[kGridBuckets];
    uint32_t item_cls[kScanPerThread], item_at[kScanPerThread];
#pragma unroll
    for (uint32_t k = 0; k < kScanPerThread; k++) item_cls[k] = 0xffffffffu;
    if (threadIdx.x < kGridBuckets) s_cls[threadIdx.x] = 0;
    __syncthreads();
    uint32_t const tiles = G.tiles_w * G.tiles_h;
    uint32_t const base = vb * kScanCells + threadIdx.x * kScanPerThread;
    uint32_t c[kScanPerThread], ns_prev[kScanPerThread];
    uint32_t sum = 0;
#pragma unroll
    for (uint32_t k = 0; k < kScanPerThread; k++)
    {
        c[k] = base + k < tiles ? G.tile_cnt[base + k] : 0u;
        // last frame's tile times, fetched with the counts: their latency hides behind the scan's barriers
        ns_prev[k] = (G.order != nullptr && base + k < tiles) ? G.tile_time[base + k] : 0u;
        sum += c[k];
    }
    uint32_t const carry = G.block_sums[vb];
    uint32_t total;
    uint32_t run = carry + block_exclusive_scan(sum, s_warp, total);
#pragma unroll
    for (uint32_t k = 0; k < kScanPerThread; k++)
        if (base + k < tiles)
        {
            G.tile_off[base + k] = run;
            G.cursor[base + k] = 0;
            G.tile_cnt[base + k] = 0;   // invariant: all counts are zero between builds (no per-frame memset)
            run += c[k];
            // Heaviest-first work list of the render launch.  A tile's cost is what it took in the PREVIOUS frame
            // (tile_time: ns, written by the render kernel, read and cleared here) -- list length is a poor
            // predictor: the slowest tiles of the benchmark frame have 8-entry lists and expensive shadow rays.
            // Tiles are cut into 2 / 4 / 8 items of whole pixel columns so that no item runs much longer than
            // ~80 us, and sorted into eight classes.  Without history (first frame, other kernels) the list length
            // decides.  Only the schedule depends on this, never a pixel.
            // entry = tile | first column << 24 | columns << 28 (0 columns = whole tile)
            if (G.order != nullptr)
            {
                uint32_t const ns = ns_prev[k];
                G.tile_time[base + k] = 0;
                if (grid_row_owned(G, static_cast<int>((base + k) / G.tiles_w)))
                {
                    bool const bvh = c[k] + G.flags->n_global > kSppMaxCand;   // too long to stage: every sample walks the BVH (~1 ms)
                    uint32_t parts, cls;
                    if (ns != 0u)
                    {
                        parts = ns >= 480000u ? 8u : ns >= 240000u ? 4u : ns >= 120000u ? 2u : 1u;
                        // eight classes, started in this order: the finer the classes, the shorter the items that are
                        // still running when the list runs dry
                        cls = ns >= 240000u ? 0u : ns >= 170000u ? 1u : ns >= 120000u ? 2u : ns >= 85000u ? 3u
                            : ns >= 60000u ? 4u : ns >= 40000u ? 5u : ns >= 12000u ? 6u : 7u;
                    }
                    else
                    {
                        parts = bvh ? 8u : 1u;
                        cls = bvh ? 0u : c[k] > 48u ? 3u : c[k] > 0u ? 5u : 7u;
                    }
                    // slot within this block's share of the class list (shared-memory atomic; the block then takes
                    // its share with ONE global atomic per class: 32 400 global atomics on four counters cost 40 us)
                    item_cls[k] = cls | (parts << 8);
                    item_at[k] = atomicAdd(&s_cls[cls], parts);
                }
            }
        }
    if (G.order != nullptr)     // grid-uniform
    {
        __syncthreads();
        if (threadIdx.x < kGridBuckets && s_cls[threadIdx.x] != 0u)
            s_cls_base[threadIdx.x] = atomicAdd(&G.bucket_cnt[threadIdx.x], s_cls[threadIdx.x]);
        __syncthreads();
#pragma unroll
        for (uint32_t k = 0; k < kScanPerThread; k++)
        {
            if (item_cls[k] == 0xffffffffu) continue;
            uint32_t const cls = item_cls[k] & 0xffu, parts = item_cls[k] >> 8, cols = 8u / parts;
            uint32_t const at = cls * G.order_stride + s_cls_base[cls] + item_at[k];
            for (uint32_t part = 0; part < parts; part++)
                G.order[at + part] = (base + k) | (parts > 1u ? ((part * cols) << 24) | (cols << 28) : 0u);
        }
    }
}

__device__ __forceinline__ void grid_phase_fill(GridBuild const& G, uint32_t vb)
{
    uint32_t const i = vb * blockDim.x + threadIdx.x;
    if (G.flags->ok == 0) return;      // grid-uniform
    float4 box = make_float4(1.f, 1.f, 0.f, 0.f);
    float4 edge[3] = {make_float4(0.f, 0.f, 1.f, 0.f), make_float4(0.f, 0.f, 1.f, 0.f), make_float4(0.f, 0.f, 1.f, 0.f)};
    if (i < G.scene.n_tris)
    {
        box = reinterpret_cast<float4 const*>(G.bbox)[i];
#pragma unroll
        for (int k = 0; k < 3; k++) edge[k] = reinterpret_cast<float4 const*>(G.edges)[3 * i + k];
    }
    int tx0 = 0, tx1 = -1, ty0 = 0, ty1 = -1;
    bool const binned = box.x <= box.z && grid_tile_range(G, box, tx0, tx1, ty0, ty1);
    // the same walk as grid_count_kernel (same boxes, same edge functions, same decisions)
    grid_walk_warp(G, binned, i, box, edge, tx0, tx1, ty0, ty1, [&](uint32_t tile, uint32_t slot, float4) {
        G.items[G.tile_off[tile] + atomicAdd(&G.cursor[tile], 1u)] = slot;
    });
}

__global__ void __launch_bounds__(256) grid_count_kernel(const __grid_constant__ GridBuild G) { grid_phase_count(G, blockIdx.x); }
__global__ void __launch_bounds__(kScanBlock) grid_block_sums_kernel(const __grid_constant__ GridBuild G) { grid_phase_block_sums(G, blockIdx.x, gridDim.x); }
__global__ void __launch_bounds__(kScanBlock) grid_scan_blocks_kernel(const __grid_constant__ GridBuild G) { grid_phase_scan_blocks(G, blockIdx.x); }
__global__ void __launch_bounds__(256) grid_fill_kernel(const __grid_constant__ GridBuild G) { grid_phase_fill(G, blockIdx.x); }

// Light grid only: every cell's list ordered by the triangles' nearest depth edges[3*slot].w, ascending (ties by
// slot, so the build is deterministic), one thread per cell, insertion sort (lists average about one entry).
__global__ void __launch_bounds__(256) sort_cells_by_depth_kernel(const __grid_constant__ GridBuild G)
{
    uint32_t const cell = blockIdx.x * blockDim.x + threadIdx.x;
    if (cell >= G.tiles_w * G.tiles_h || G.flags->ok == 0) return;
    uint32_t const b = G.tile_off[cell], e = G.tile_off[cell + 1];
    float4 const* edges = reinterpret_cast<float4 const*>(G.edges);
    for (uint32_t i = b + 1; i < e; i++)
    {
        uint32_t const s = G.items[i];
        float const d = edges[3 * s].w;
        uint32_t j = i;
        while (j > b)
        {
            uint32_t const p = G.items[j - 1];
            float const dp = edges[3 * p].w;
            if (dp < d || (dp == d && p < s)) break;
            G.items[j] = p;
            j--;
        }
        G.items[j] = s;
    }
}

// tri_n[i] = normalize(e1[i] x e2[i]) with the spec's operation sequence (docs/RENDER_SPEC.md section 4): the
// bits every shaded sample used to recompute.  Degenerate triangles give the same NaN/inf as before.
__global__ void __launch_bounds__(256) tri_normals_kernel(float4 const* __restrict__ e1, float4 const* __restrict__ e2,
                                                          float4* __restrict__ n, uint32_t count)
{
    uint32_t const i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i >= count) return;
    V3 const nr = cross3(v3(e1[i]), v3(e2[i]));
    float const nl = __fsqrt_rn(dot3(nr, nr));
    float const inl = __fdiv_rn(1.0f, nl);
    n[i] = make_float4(mul(nr.x, inl), mul(nr.y, inl), mul(nr.z, inl), 0.f);
}

// ---------------------------------------------------------------------------
// BVH traversal
// ---------------------------------------------------------------------------
constexpr int kBvhThreads = 256;  // 8 warps; each warp owns an 8x4 pixel tile
constexpr int kBvhBlockW = 32;
constexpr int kBvhBlockH = 8;

struct RayBox
{
    V3 idir; // 1/d with |d| clamped away from 0
    V3 ood;  // o * idir
};

__device__ __forceinline__ float safe_inv(float d)
{
    float const eps = 1e-30f;
    float a = fabsf(d) > eps ? d : copysignf(eps, d);
    return 1.0f / a;
}

// Culling only (docs/RENDER_SPEC.md section 6): boxes are padded at build time and
// `t_far` carries a relative slack, so rounding here can only visit too much.
//
// `hint` (kNoHint = none) is the slot of a triangle a neighbouring ray hit: it is tested BEFORE the
// traversal.  Closest hit: the traversal then starts with a tight t_far and culls every subtree
// behind it; any hit: a still-valid occluder ends the query without touching the BVH.  The result
// is unchanged -- the closest-hit winner is the lexicographic minimum of (t, id) whatever the test
// order, occlusion is a boolean -- only the number of nodes visited drops.  On return best.slot
// holds the occluder's slot when ANY found one (the next ray's hint).
constexpr uint32_t kNoHint = 0xffffffffu;

template <bool ANY, bool STATS>
__device__ __forceinline__ void bvh_trace(DeviceSceneView const& sc, V3 o, V3 d, float t_min,
                                          Hit& best, bool& occluded,
                                          unsigned long long& n_box, unsigned long long& n_tri,
                                          uint32_t hint = kNoHint)
{
    float4 const* __restrict__ nodes = reinterpret_cast<float4 const*>(sc.nodes);
    float4 const* __restrict__ tv0 = reinterpret_cast<float4 const*>(sc.tri_v0);
    float4 const* __restrict__ te1 = reinterpret_cast<float4 const*>(sc.tri_e1);
    float4 const* __restrict__ te2 = reinterpret_cast<float4 const*>(sc.tri_e2);

    // Slab test of BOTH children at once with packed f32x2 FMAs (SASS FFMA2: one
    // issue slot, two lanes of work).  The near/far plane of an axis is picked by
    // the sign of the direction WITHOUT min/max: with idp = max(1/d,0) and
    // idn = min(1/d,0) exactly one is non-zero, so
    //     t_near = min_plane*idp + max_plane*idn - o/d ,  t_far = max_plane*idp + min_plane*idn - o/d
    // which moves 12 FMNMX per node from the half-rate ALU pipe to 12 FFMA2.
    float const idx = safe_inv(d.x), idy = safe_inv(d.y), idz = safe_inv(d.z);
    F2 const idpx = pack2(fmaxf(idx, 0.f)), idnx = pack2(fminf(idx, 0.f)), noodx = pack2(-(o.x * idx));
    F2 const idpy = pack2(fmaxf(idy, 0.f)), idny = pack2(fminf(idy, 0.f)), noody = pack2(-(o.y * idy));
    F2 const idpz = pack2(fmaxf(idz, 0.f)), idnz = pack2(fminf(idz, 0.f)), noodz = pack2(-(o.z * idz));
    float const slack = 1.000244140625f; // 1 + 2^-12
    float t_far = ANY ? slack : pos_inf();
    float const t_near = fmaxf(t_min, 0.0f);

    if (!STATS && hint != kNoHint)
    {
        float4 v0 = __ldg(tv0 + hint), e1 = __ldg(te1 + hint), e2 = __ldg(te2 + hint);
        float ts, adet;
        if (mt_inside(o, d, v0, e1, e2, ts, adet))
        {
            if (ANY)
            {
                float t = __fdiv_rn(ts, adet);
                if (t > 0.0f && t < 1.0f) { occluded = true; best.slot = hint; return; }
            }
            else
            {
                closest_update(best, ts, adet, __float_as_int(v0.w), hint, t_min);
                t_far = best.t * slack; // +inf * slack = +inf when the hint was rejected
            }
        }
    }

    int stack[kBvhStackDepth];
    int sp = 0;
    int node = 0; // root is always an inner node

    while (true)
    {
        while (node >= 0)
        {
            // one 64-byte node: a single widening multiply-add for the address
            float4 const* const np = reinterpret_cast<float4 const*>(
                reinterpret_cast<char const*>(nodes) + static_cast<size_t>(static_cast<uint32_t>(node)) * 64u);
            float4 n0 = __ldg(np + 0);
            float4 n1 = __ldg(np + 1);
            float4 n2 = __ldg(np + 2);
            float4 n3 = __ldg(np + 3);
            if (STATS) n_box += 2;

            // n = (c0.min, c1.min, c0.max, c1.max) per axis: .lo pair = mins, .hi pair = maxes
            F2 const mnx = pack2(n0.x, n0.y), mxx = pack2(n0.z, n0.w);
            F2 const mny = pack2(n1.x, n1.y), mxy = pack2(n1.z, n1.w);
            F2 const mnz = pack2(n2.x, n2.y), mxz = pack2(n2.z, n2.w);
            float2 const nx = unpack2(fma2(mnx, idpx, fma2(mxx, idnx, noodx)));
            float2 const fx = unpack2(fma2(mxx, idpx, fma2(mnx, idnx, noodx)));
            float2 const ny = unpack2(fma2(mny, idpy, fma2(mxy, idny, noody)));
            float2 const fy = unpack2(fma2(mxy, idpy, fma2(mny, idny, noody)));
            float2 const nz = unpack2(fma2(mnz, idpz, fma2(mxz, idnz, noodz)));
            float2 const fz = unpack2(fma2(mxz, idpz, fma2(mnz, idnz, noodz)));

            float c0min = fmaxf(fmaxf(nx.x, ny.x), fmaxf(nz.x, t_near));
            float c0max = fminf(fminf(fx.x, fy.x), fminf(fz.x, t_far));
            float c1min = fmaxf(fmaxf(nx.y, ny.y), fmaxf(nz.y, t_near));
            float c1max = fminf(fminf(fx.y, fy.y), fminf(fz.y, t_far));

            // No slack on the entry side: the build-time padding (2^-18 of the scene
            // extent, absolute) exceeds the slab arithmetic's rounding by > 10x.
            bool hit0 = c0max >= c0min;
            bool hit1 = c1max >= c1min;
            int ref0 = __float_as_int(n3.x), ref1 = __float_as_int(n3.y);

            if (!hit0 && !hit1)
            {
                if (sp == 0) return;
                node = stack[--sp];
            }
            else if (hit0 && hit1)
            {
                bool swap = c1min < c0min;
                node = swap ? ref1 : ref0;
                stack[sp++] = swap ? ref0 : ref1;
            }
            else
            {
                node = hit0 ? ref0 : ref1;
            }
        }

        // leaf
        {
            uint32_t enc = static_cast<uint32_t>(~node);
            uint32_t first = enc >> kLeafCountBits;
            uint32_t count = (enc & ((1u << kLeafCountBits) - 1u)) + 1u;
            if (STATS) n_tri += count;
            for (uint32_t i = first; i < first + count; i++)
            {
                float4 v0 = __ldg(tv0 + i), e1 = __ldg(te1 + i), e2 = __ldg(te2 + i);
                float ts, adet;
                if (mt_inside(o, d, v0, e1, e2, ts, adet))
                {
                    if (ANY)
                    {
                        float t = __fdiv_rn(ts, adet);
                        if (t > 0.0f && t < 1.0f) { occluded = true; best.slot = i; return; }
                    }
                    else
                    {
                        closest_update(best, ts, adet, __float_as_int(v0.w), i, t_min);
                        t_far = best.t * slack;
                    }
                }
            }
        }
        if (sp == 0) return;
        node = stack[--sp];
    }
}

// Shadow query through the light grid (scene.h LightGridView).  Returns false when the grid cannot
// answer (shading point behind the light's projection plane, non-finite projection): the caller walks
// the BVH.  The occluder set is exact: a point of a binned triangle between S and the light projects
// to S's own cell (boxes dilated for every rounding involved), triangles that could not be binned are
// in the global list, triangles wholly behind the plane cannot lie between a point in front and the light.
// Culling arithmetic below (projection, edge functions) never enters the image: it uses explicit FMAs and an
// approximate reciprocal; the grid's dilation budgets ~34 ulps of relative error for this projection.
__device__ __forceinline__ float rcp_approx(float x)
{
    float r;
    asm("rcp.approx.ftz.f32 %0, %1;" : "=f"(r) : "f"(x));
    return r;
}
__device__ __forceinline__ bool edge_pass(float4 e, float qx, float qy)
{
    return __fmaf_rn(e.x, qx, __fmaf_rn(e.y, qy, e.z)) >= 0.0f;
}

template <bool STATS>
__device__ __forceinline__ bool lgrid_shadow(RenderLaunch const& L, V3 so, V3 sd, bool& occluded, uint32_t& hint,
                                             uint32_t n_global, unsigned long long& n_tri)
{
    LightGridView const& G = L.lgrid;
    float const px = so.x - L.light.position[0], py = so.y - L.light.position[1], pz = so.z - L.light.position[2];
    float const cz = __fmaf_rn(G.minv[6], px, __fmaf_rn(G.minv[7], py, G.minv[8] * pz));
    if (!(cz > 0.0f)) return false;
    float const iz = rcp_approx(cz);
    float const cx = __fmaf_rn(G.minv[0], px, __fmaf_rn(G.minv[1], py, G.minv[2] * pz));
    float const cy = __fmaf_rn(G.minv[3], px, __fmaf_rn(G.minv[4], py, G.minv[5] * pz));
    float const gx = __fmaf_rn(cx * iz, G.kx, G.cx);
    float const gy = __fmaf_rn(-(cy * iz), G.ky, G.cy);
    if (!(fabsf(gx) < 1e9f) || !(fabsf(gy) < 1e9f)) return false;
    int const ix = min(max(static_cast<int>(floorf(gx)), 0), static_cast<int>(G.grid.tiles_w) - 1);
    int const iy = min(max(static_cast<int>(floorf(gy)), 0), static_cast<int>(G.grid.tiles_h) - 1);
    uint32_t const cell = static_cast<uint32_t>(iy) * G.grid.tiles_w + static_cast<uint32_t>(ix);
    float4 const* __restrict__ tv0 = reinterpret_cast<float4 const*>(L.scene.tri_v0);
    float4 const* __restrict__ te1 = reinterpret_cast<float4 const*>(L.scene.tri_e1);
    float4 const* __restrict__ te2 = reinterpret_cast<float4 const*>(L.scene.tri_e2);
    auto test = [&](uint32_t slot) -> bool {
        float4 const v0 = __ldg(tv0 + slot), e1 = __ldg(te1 + slot), e2 = __ldg(te2 + slot);
        float ts, adet;
        if (mt_inside(so, sd, v0, e1, e2, ts, adet))
        {
            float const t = __fdiv_rn(ts, adet);
            if (t > 0.0f && t < 1.0f) { occluded = true; hint = slot; return true; }
        }
        return false;
    };
    if (hint != kNoHint)
    {
        if (STATS) n_tri++;
        if (test(hint)) return true;
    }
    for (uint32_t k = 0; k < n_global; k++)
    {
        if (STATS) n_tri++;
        if (test(__ldg(G.grid.global + k))) return true;
    }
    uint32_t const b = __ldg(G.grid.tile_off + cell), e = __ldg(G.grid.tile_off + cell + 1);
    float4 const* __restrict__ edges = reinterpret_cast<float4 const*>(G.grid.edges);
    // The cell's list is sorted by the triangles' nearest depth (sort_cells_by_depth_kernel): a triangle can only
    // lie between the shading point and the light if it reaches nearer to the light than the point, so the walk
    // ends at the first entry beyond it.
    // A cell seen at a grazing angle from the light can list hundreds of triangles (measured: tiles whose shadow
    // rays spent 13K instructions each in this loop); after max_walk entries the BVH decides the rest (exact).
    uint32_t const e_walk = min(e, b + G.max_walk);
    for (uint32_t k = b; k < e; k++)
    {
        uint32_t const slot = __ldg(G.grid.items + k);
        // One round trip for everything the entry may need (edge functions and the triangle itself) instead of up
        // to four dependent ones: with ~1 entry per cell, most entries visited are the ones that cover the point.
        float4 const e0 = __ldg(edges + 3 * slot), ea = __ldg(edges + 3 * slot + 1), eb = __ldg(edges + 3 * slot + 2);
        float4 const v0 = __ldg(tv0 + slot), e1 = __ldg(te1 + slot), e2 = __ldg(te2 + slot);
        if (!(e0.w <= cz)) break;
        if (k >= e_walk) return false;
        // the projected triangle must cover the query point
        if (edge_pass(e0, gx, gy) & edge_pass(ea, gx, gy) & edge_pass(eb, gx, gy))
        {
            if (STATS) n_tri++;
            float ts, adet;
            if (mt_inside(so, sd, v0, e1, e2, ts, adet))
            {
                float const t = __fdiv_rn(ts, adet);
                if (t > 0.0f && t < 1.0f) { occluded = true; hint = slot; return true; }
            }
        }
    }
    return true;
}

// MINB = minimum resident blocks per SM the register allocation must allow.
constexpr uint32_t kPixGridMaxList = 160;  // thread-per-pixel kernel: longer tile lists walk the BVH

template <bool STATS, int MINB>
__global__ void __launch_bounds__(kBvhThreads, MINB) render_bvh_kernel(const __grid_constant__ RenderLaunch L)
{
    uint32_t const W = L.params.width;
    uint32_t const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t const x = blockIdx.x * kBvhBlockW + (warp & 3u) * 8u + (lane & 7u);
    uint32_t const y = L.row_origin + blockIdx.y * L.row_block_step + (warp >> 2) * 4u + (lane >> 3);
    bool const valid = x < W && y < L.params.y1;
    unsigned long long const t_block = L.d_trace ? trace_now() : 0ull;
    // a warp's 8x4 pixels are four 96-byte row segments of the frame: staged here, written as 128-bit stores
    __shared__ __align__(16) float s_out[kBvhThreads / 32][4][24];
    V3 out_rgb = V3{0.f, 0.f, 0.f};

    unsigned long long n_primary = 0, n_shadow = 0, n_box = 0, n_tri = 0;
    if (valid)
    {
        // Screen grid: the 32 pixels of a warp lie in ONE 8x8 tile, so every lane walks the same candidate
        // list (uniform addresses: broadcast loads) and filters it against its own pixel.
        bool use_grid = false;
        uint32_t g_off = 0, g_cnt = 0, g_glob = 0;
        if (L.grid.enabled)
        {
            GridFlags const fl = *L.grid.flags;
            uint32_t const tile = (y / kGridTile) * L.grid.tiles_w + x / kGridTile;
            g_off = __ldg(L.grid.tile_off + tile);
            g_cnt = __ldg(L.grid.tile_off + tile + 1) - g_off;
            g_glob = fl.n_global;
            use_grid = fl.ok != 0 && g_cnt <= kPixGridMaxList;
        }
        bool lgrid_ok = false;
        uint32_t lgrid_global = 0;
        if (L.lgrid.grid.enabled)
        {
            GridFlags const lf = *L.lgrid.grid.flags;
            lgrid_ok = lf.ok != 0;
            lgrid_global = lf.n_global;
        }
        V3 const o = v3(L.camera.origin);
        V3 acc = V3{0.f, 0.f, 0.f};
        Hit first_hit = no_hit();
        for (uint32_t s = 0; s < L.params.spp; s++)
        {
            V3 d = primary_dir(L, x, y, s);
            Hit best = no_hit();
            bool occ = false;
            n_primary++;
            if (use_grid)
            {
                float const px0 = static_cast<float>(x), py0 = static_cast<float>(y);
                float const qx = px0 + 0.5f, qy = py0 + 0.5f;
                float4 const* __restrict__ edges = reinterpret_cast<float4 const*>(L.grid.edges);
                for (uint32_t k = 0; k < g_cnt + g_glob; k++)
                {
                    bool const glob = k >= g_cnt;
                    uint32_t const slot = glob ? __ldg(L.grid.global + (k - g_cnt)) : __ldg(L.grid.items + g_off + k);
                    if (!glob)
                    {
                        // box and edge functions in one round trip (the box alone rejects few entries of a tile's list)
                        float4 const b = __ldg(reinterpret_cast<float4 const*>(L.grid.bbox) + slot);
                        float4 const e0 = __ldg(edges + 3 * slot), e1 = __ldg(edges + 3 * slot + 1), e2 = __ldg(edges + 3 * slot + 2);
                        if (!(b.z >= px0 && b.x <= px0 + 1.0f && b.w >= py0 && b.y <= py0 + 1.0f)) continue;
                        if (!(e0.x * qx + e0.y * qy + e0.z >= 0.0f) || !(e1.x * qx + e1.y * qy + e1.z >= 0.0f)
                            || !(e2.x * qx + e2.y * qy + e2.z >= 0.0f))
                            continue;
                    }
                    if (STATS) n_tri++;
                    float4 const v0 = __ldg(reinterpret_cast<float4 const*>(L.scene.tri_v0) + slot);
                    float4 const e1 = __ldg(reinterpret_cast<float4 const*>(L.scene.tri_e1) + slot);
                    float4 const e2 = __ldg(reinterpret_cast<float4 const*>(L.scene.tri_e2) + slot);
                    float ts, adet;
                    if (mt_inside(o, d, v0, e1, e2, ts, adet))
                        closest_update(best, ts, adet, __float_as_int(v0.w), slot, L.params.t_min);
                }
            }
            else
                bvh_trace<false, STATS>(L.scene, o, d, L.params.t_min, best, occ, n_box, n_tri);
            if (s == 0) first_hit = best;
            V3 rgb;
            if (best.id < 0)
            {
                rgb = v3(L.params.background);
            }
            else
            {
                ShadeOut sh = shade_hit(L, o, d, best);
                rgb = sh.rgb;
                if (sh.need_shadow)
                {
                    n_shadow++;
                    Hit dummy = no_hit();
                    dummy.slot = kNoHint;
                    if (!(lgrid_ok && lgrid_shadow<STATS>(L, sh.so, sh.sd, occ, dummy.slot, lgrid_global, n_tri)))
                        bvh_trace<true, STATS>(L.scene, sh.so, sh.sd, 0.0f, dummy, occ, n_box, n_tri);
                    if (occ) rgb = V3{0.f, 0.f, 0.f};
                }
            }
            acc.x = add(acc.x, rgb.x);
            acc.y = add(acc.y, rgb.y);
            acc.z = add(acc.z, rgb.z);
        }
        out_rgb = V3{mul(acc.x, L.inv_spp), mul(acc.y, L.inv_spp), mul(acc.z, L.inv_spp)};
        size_t const out = static_cast<size_t>(y * W + x) - static_cast<size_t>(L.params.buf_y0) * W;
        if (L.d_hit_id) L.d_hit_id[out] = first_hit.id;
        if (L.d_hit_t) L.d_hit_t[out] = first_hit.t;
    }
    if (L.d_rgb)
    {
        // whole, 16-byte aligned segments (every tile of a frame whose width is a multiple of 4, edges aside):
        // 24 lanes store one float4 each; otherwise every valid lane stores its own three floats
        uint32_t const lx = lane & 7u, ly = lane >> 3;
        uint32_t const x0 = x - lx, y0 = y - ly;
        float* const seg = L.d_rgb + (static_cast<size_t>(y0 * W + x0) - static_cast<size_t>(L.params.buf_y0) * W) * 3;
        bool const vec = __all_sync(0xffffffffu, valid) && (W & 3u) == 0u && (reinterpret_cast<uintptr_t>(seg) & 15u) == 0;
        if (vec)
        {
            float* const mine = &s_out[warp][ly][lx * 3u];
            mine[0] = out_rgb.x; mine[1] = out_rgb.y; mine[2] = out_rgb.z;
            __syncwarp();
            if (lane < 24u)
            {
                uint32_t const r = lane / 6u, v = lane - r * 6u;
                reinterpret_cast<float4*>(seg + static_cast<size_t>(r) * W * 3u)[v] =
                    reinterpret_cast<float4 const*>(s_out[warp][r])[v];
            }
        }
        else if (valid)
        {
            float* const px = seg + (static_cast<size_t>(ly) * W + lx) * 3u;
            px[0] = out_rgb.x; px[1] = out_rgb.y; px[2] = out_rgb.z;
        }
    }
    trace_block(L, t_block);
    flush_counters(L, n_primary, n_shadow, n_box, n_tri);
}

// Sample-parallel variant for spp >= 32/PIX.  A warp works on PIX pixels at a
// time and spreads each pixel's samples over SPL = 32/PIX lanes, so the lanes
// of a warp trace rays through (nearly) the same pixel: they walk the same BVH
// path and diverge far less than 32 different pixels do, and a block is a
// 1/32-th the work of the pixel-per-thread kernel (short tail).  Radiance is
// still summed in sample order: each pass parks the SPL sample values of a
// pixel in shared memory and ONE lane per (pixel, channel) adds them
// sequentially onto its running sum, so the result is bit-identical to the
// pixel-per-thread kernel and the oracle (docs/RENDER_SPEC.md section 5).
constexpr int kSppTileW = 8; // block = 8 warps; warp w owns row w of an 8x8 pixel tile

// Moller-Trumbore against a staged candidate whose ray-ORIGIN-dependent terms are precomputed: all primary
// rays leave the camera origin o, so tv = o - v0, q = tv x e1 and tsr = e2 . q are per-triangle constants of
// the frame.  They are produced by the same operations on the same operands as in mt_inside, hence the
// same bits; what is left per ray is p = d x e2, det = e1 . p, us = tv . p, vs = d . q: 15 instead of 33
// FP32 instructions.
__device__ __forceinline__ bool mt_inside_pre(V3 d, float4 e1_tsr, float4 e2_id, float4 tv4, float4 q4,
                                              float& ts, float& adet)
{
    V3 const E1 = v3(e1_tsr), E2 = v3(e2_id), tv = v3(tv4), q = v3(q4);
    V3 const p = cross3(d, E2);
    float const det = dot3(E1, p);
    float us = dot3(tv, p);
    float vs = dot3(d, q);
    uint32_t const sgn = __float_as_uint(det) & 0x80000000u;
    us = __uint_as_float(__float_as_uint(us) ^ sgn);
    vs = __uint_as_float(__float_as_uint(vs) ^ sgn);
    ts = __uint_as_float(__float_as_uint(e1_tsr.w) ^ sgn);
    adet = fabsf(det);
    return (us >= 0.0f) & (vs >= 0.0f) & (add(us, vs) <= adet);
}

template <bool STATS, int PIX, int MINB = 4, bool PERSIST = false>
__global__ void __launch_bounds__(256, MINB) render_bvh_spp_kernel(const __grid_constant__ RenderLaunch L)
{
    constexpr uint32_t SPL = 32u / PIX; // sample lanes per pixel
    if (L.d_trace != nullptr && threadIdx.x == 0) atomicMax(L.d_trace + 12, ~trace_now());   // earliest block entry
    static_assert(SPL >= 4, "one lane per (pixel, channel) does the ordered sum; 128-bit reads of the sample values");
    // sample values of the current pass, channel-major so that the summing lane reads them as float4
    __shared__ __align__(16) float s_val[8][3][32];
    // finished pixels of the warp's row (8 pixels x RGB): written to the frame as 128-bit stores
    __shared__ __align__(16) float s_out[8][24];
    // Primary-ray grid (PIX == 1 only): the block IS one 8x8 grid tile.  The tile's candidate triangles
    // (plus the frame's global ones) are staged in shared memory once -- edges, the origin-dependent
    // Moller-Trumbore terms, unit normal and albedo, so that neither the test nor the shading of a sample
    // touches global memory; every warp filters them against its current pixel (lanes test different
    // candidates, ballot-compacted), and every sample of that pixel then tests the few survivors.
    constexpr uint32_t kCand = PIX == 1 ? kSppMaxCand : 1;
    static_assert(kCand <= 256, "s_pixlist holds candidate indices in one byte; one staging thread per candidate");
    __shared__ float4 s_ce1[kCand], s_ce2[kCand], s_ctv[kCand], s_cq[kCand], s_cn[kCand], s_calb[kCand];
    __shared__ float4 s_cbox[kCand], s_cedge[3][kCand];
    __shared__ uint8_t s_pixlist[PIX == 1 ? 8 : 1][kCand];
    __shared__ float s_key[kCand];      // depth keys of the candidates while they are being ranked
    V3 const o = v3(L.camera.origin);
    uint32_t const W = L.params.width;
    uint32_t const warp = threadIdx.x >> 5, lane = threadIdx.x & 31;
    uint32_t const p = lane / SPL, sl = lane % SPL;
    uint32_t const spp = L.params.spp;
    V3 const bg = v3(L.params.background);

    unsigned long long n_primary = 0, n_shadow = 0, n_box = 0, n_tri = 0;
    // triangle this lane's previous primary / shadow ray ended on: the next ray of the same (or the
    // neighbouring) pixel almost always ends on it too
    uint32_t hint_hit = kNoHint, hint_occ = kNoHint;
    bool const use_hints = L.hints != 0;
    bool lgrid_ok = false;
    uint32_t lgrid_global = 0;
    if (L.lgrid.grid.enabled)
    {
        GridFlags const lf = *L.lgrid.grid.flags;
        lgrid_ok = lf.ok != 0;
        lgrid_global = lf.n_global;
    }

    // Work items.  Plain launch: the block renders the 8x8 pixel tile (blockIdx.x, owned row block blockIdx.y)
    // and is done.  Ordered launch (RenderLaunch::ordered, PIX == 1): the blocks are PERSISTENT (one per
    // resident slot) and pull items from the list the grid build left in GridView::order -- heaviest class
    // first, costly tiles cut into 2 / 4 / 8 items of whole pixel columns -- through an atomic cursor, so the
    // launch neither pays a block start per tile nor drains on a millisecond-long tile (the longest tile of the
    // benchmark frame takes 1.1 ms: more than a whole frame should take on a rank of eight).
    constexpr bool persistent = PIX == 1 && PERSIST;      // compile-time: the plain kernel carries none of this
    uint32_t n_items = 0;
    if (persistent)
        for (uint32_t c = 0; c < kGridBuckets; c++) n_items += L.grid.bucket_cnt[c];
    __shared__ uint32_t s_item;
    // thread 0 draws the NEXT item's index while the current item is being rendered (the atomic's round trip to
    // L2 hides behind the work) and publishes it at the barrier that ends the item
    uint32_t next_item = 0;
    if (persistent && threadIdx.x == 0) s_item = atomicAdd(L.d_work, 1u);
    for (;;)
    {
    uint32_t bx = blockIdx.x, row0 = L.row_origin + blockIdx.y * L.row_block_step;
    uint32_t g_begin = 0, g_end = kSppTileW / PIX;      // pixel columns (groups) of the tile this item covers
    if (persistent)
    {
        __syncthreads();                 // s_item published; every warp has left the previous item's shared arrays
        uint32_t i = s_item, cls = 0;
        if (i >= n_items) break;         // block-uniform
        if (threadIdx.x == 0) next_item = atomicAdd(L.d_work, 1u);
        for (; cls < kGridBuckets - 1u; cls++)
        {
            uint32_t const n = L.grid.bucket_cnt[cls];
            if (i < n) break;
            i -= n;
        }
        uint32_t const entry = L.grid.order[cls * L.grid.order_stride + i];
        uint32_t const t = entry & 0xffffffu, cols = entry >> 28;     // cols 0: whole tile; else `cols` columns from bits 24..27
        uint32_t const tyo = t / L.grid.tiles_w;
        bx = t - tyo * L.grid.tiles_w;
        row0 = tyo * kGridTile;
        if (cols) { g_begin = (entry >> 24) & 15u; g_end = g_begin + cols; }
    }
    bool const timed_item = persistent && L.grid.enabled && L.grid.tile_time != nullptr;
    unsigned long long const t_block = (L.d_trace || timed_item) ? trace_now() : 0ull;
    unsigned long long const box0 = n_box, tri0 = n_tri;
    uint32_t const shadow0 = n_shadow;
    uint32_t n_cand = 0, trace_tile = 0, trace_cnt = 0;
    bool use_grid = false;
    if (PIX == 1 && L.grid.enabled)
    {
        GridFlags const fl = *L.grid.flags;
        uint32_t const ty = row0 / kGridTile;
        uint32_t const tile = ty * L.grid.tiles_w + bx;
        uint32_t const off = L.grid.tile_off[tile], cnt = L.grid.tile_off[tile + 1] - off;
        trace_tile = tile; trace_cnt = cnt;
        use_grid = fl.ok != 0 && cnt + fl.n_global <= kCand && bx < L.grid.tiles_w && ty < L.grid.tiles_h;
        if (use_grid)   // block-uniform
        {
            n_cand = cnt + fl.n_global;
            // The candidates are staged NEAREST FIRST (rank sort on the depth key the grid build left in
            // edges[3*slot].w: nearest vertex depth along `forward` minus the projection slack; global triangles
            // -inf): a sample stops walking its pixel's list at the first candidate that lies wholly behind the hit
            // it already has, which skips the hidden layers (the benchmark frame: 1.8 layers per covered pixel).
            bool const mine = threadIdx.x < n_cand;
            bool const glob = threadIdx.x >= cnt;
            uint32_t slot = 0;
            float4 edge0 = make_float4(0.f, 0.f, 1.f, -pos_inf());
            if (mine)
            {
                slot = glob ? L.grid.global[threadIdx.x - cnt] : L.grid.items[off + threadIdx.x];
                if (!glob) edge0 = __ldg(reinterpret_cast<float4 const*>(L.grid.edges) + 3 * slot);
                s_key[threadIdx.x] = edge0.w;
            }
            __syncthreads();
            if (mine)
            {
                uint32_t at = 0;
                for (uint32_t k = 0; k < n_cand; k++)
                {
                    float const other = s_key[k];      // same address in every lane: broadcast
                    at += (other < edge0.w || (other == edge0.w && k < threadIdx.x)) ? 1u : 0u;
                }
                float4 const v0 = __ldg(reinterpret_cast<float4 const*>(L.scene.tri_v0) + slot);
                float4 const e1 = __ldg(reinterpret_cast<float4 const*>(L.scene.tri_e1) + slot);
                float4 const e2 = __ldg(reinterpret_cast<float4 const*>(L.scene.tri_e2) + slot);
                V3 const tv = V3{sub(o.x, v0.x), sub(o.y, v0.y), sub(o.z, v0.z)};
                V3 const q = cross3(tv, v3(e1));
                s_ce1[at] = make_float4(e1.x, e1.y, e1.z, dot3(v3(e2), q));
                s_ce2[at] = make_float4(e2.x, e2.y, e2.z, v0.w);          // .w = bits(triangle id)
                s_ctv[at] = make_float4(tv.x, tv.y, tv.z, 0.f);
                s_cq[at] = make_float4(q.x, q.y, q.z, edge0.w);           // .w = depth key
                s_cn[at] = __ldg(reinterpret_cast<float4 const*>(L.scene.tri_n) + slot);
                s_calb[at] = __ldg(reinterpret_cast<float4 const*>(L.scene.albedo) + __float_as_int(v0.w));
                s_cbox[at] = glob ? make_float4(-pos_inf(), -pos_inf(), pos_inf(), pos_inf())
                                  : __ldg(reinterpret_cast<float4 const*>(L.grid.bbox) + slot);
                s_cedge[0][at] = edge0;
#pragma unroll
                for (int k = 1; k < 3; k++)
                    s_cedge[k][at] = glob ? make_float4(0.f, 0.f, 1.f, 0.f)
                                          : __ldg(reinterpret_cast<float4 const*>(L.grid.edges) + 3 * slot + k);
            }
            __syncthreads();
        }
    }

    uint32_t const y = row0 + warp;
    if (y < L.params.y1) // warp-uniform
    {
        for (uint32_t g = g_begin; g < g_end; g++)
        {
            uint32_t const x = bx * kSppTileW + g * PIX + p;
            bool const in_frame = x < W;
            float sum = 0.f; // running sum of channel `sl` of pixel p (lanes with sl < 3)
            Hit first_hit = no_hit();
            uint32_t n_pix = 0; // candidates whose screen box touches this pixel (closed intervals)
            if (PIX == 1 && use_grid)
            {
                float const px0 = static_cast<float>(x), py0 = static_cast<float>(y);
                float const qx = px0 + 0.5f, qy = py0 + 0.5f;   // the margin in c covers the pixel square
                for (uint32_t base = 0; base < n_cand; base += 32u)
                {
                    uint32_t const j = base + lane;
                    bool pass = false;
                    if (j < n_cand)
                    {
                        float4 const b = s_cbox[j];
                        pass = b.z >= px0 && b.x <= px0 + 1.0f && b.w >= py0 && b.y <= py0 + 1.0f;
#pragma unroll
                        for (int k = 0; k < 3; k++) pass = pass && edge_pass(s_cedge[k][j], qx, qy);
                    }
                    uint32_t const m = __ballot_sync(0xffffffffu, pass);
                    if (pass) s_pixlist[warp][n_pix + __popc(m & ((1u << lane) - 1u))] = static_cast<uint8_t>(j);
                    n_pix += __popc(m);
                }
                __syncwarp();
            }
            for (uint32_t s0 = 0; s0 < spp; s0 += SPL)
            {
                uint32_t const s = s0 + sl;
                V3 rgb = V3{0.f, 0.f, 0.f};
                if (in_frame && s < spp)
                {
                    float len;
                    V3 d = primary_dir(L, x, y, s, &len);
                    Hit best = no_hit();
                    bool occ = false;
                    n_primary++;
                    ShadeOut sh;
                    sh.need_shadow = false;
                    if (PIX == 1 && use_grid)
                    {
                        // A candidate whose depth key is k lies at t >= k * len along this ray.  Stop when that is
                        // beyond the hit in hand by more than 1/64 -- a margin some 10^5 ulps wide for the rounding of
                        // the two t values being compared (the spec's winner is the minimum of the COMPUTED t's).
                        float const t_scale = mul(len, 0.984375f);
                        for (uint32_t k = 0; k < n_pix; k++)
                        {
                            uint32_t const j = s_pixlist[warp][k];       // same address in every lane: broadcast
                            float4 const cq = s_cq[j];
                            if (best.t < mul(cq.w, t_scale)) break;      // nearest first: so are all the others
                            if (STATS) n_tri++;
                            float4 const e1 = s_ce1[j], e2 = s_ce2[j];
                            float ts, adet;
                            if (mt_inside_pre(d, e1, e2, s_ctv[j], cq, ts, adet))
                                closest_update(best, ts, adet, __float_as_int(e2.w), j, L.params.t_min);   // slot = list index
                        }
                        if (best.id >= 0) sh = shade_hit(L, o, d, best.t, v3(s_cn[best.slot]), v3(s_calb[best.slot]));
                    }
                    else
                    {
                        bvh_trace<false, STATS>(L.scene, o, d, L.params.t_min, best, occ, n_box, n_tri, hint_hit);
                        if (best.id >= 0)
                        {
                            sh = shade_hit(L, o, d, best);
                            if (use_hints) hint_hit = best.slot;
                        }
                    }
                    if (s == 0) first_hit = best;
                    if (best.id < 0)
                    {
                        rgb = bg;
                    }
                    else
                    {
                        rgb = sh.rgb;
                        if (sh.need_shadow)
                        {
                            n_shadow++;
                            Hit dummy = no_hit();
                            dummy.slot = hint_occ;       // in: the last occluder (tested first), out: this ray's occluder
                            if (!(lgrid_ok && lgrid_shadow<STATS>(L, sh.so, sh.sd, occ, dummy.slot, lgrid_global, n_tri)))
                                bvh_trace<true, STATS>(L.scene, sh.so, sh.sd, 0.0f, dummy, occ, n_box, n_tri, hint_occ);
                            if (occ)
                            {
                                rgb = V3{0.f, 0.f, 0.f};
                                if (use_hints) hint_occ = dummy.slot;
                            }
                        }
                    }
                }
                s_val[warp][0][lane] = rgb.x;
                s_val[warp][1][lane] = rgb.y;
                s_val[warp][2][lane] = rgb.z;
                __syncwarp();
                if (sl < 3)
                {
                    // ordered sum (docs/RENDER_SPEC.md section 5): sample values strictly in index order
                    float const* const v = &s_val[warp][sl][p * SPL];
                    uint32_t const n = min(SPL, spp - s0);
                    if (n == SPL)
                    {
#pragma unroll
                        for (uint32_t j = 0; j < SPL; j += 4)
                        {
                            float4 const q4 = *reinterpret_cast<float4 const*>(v + j);
                            sum = add(add(add(add(sum, q4.x), q4.y), q4.z), q4.w);
                        }
                    }
                    else
                        for (uint32_t j = 0; j < n; j++) sum = add(sum, v[j]);
                }
                __syncwarp();
            }
            if (sl < 3) s_out[warp][(g * PIX + p) * 3 + sl] = mul(sum, L.inv_spp);
            if (in_frame && sl == 0) // this lane traced sample 0
            {
                size_t const out = static_cast<size_t>(y * W + x) - static_cast<size_t>(L.params.buf_y0) * W;
                if (L.d_hit_id) L.d_hit_id[out] = first_hit.id;
                if (L.d_hit_t) L.d_hit_t[out] = first_hit.t;
            }
        }
        // the warp's 8 pixels x RGB = 96 contiguous bytes of the frame: six 128-bit stores in one instruction
        // when the row segment is whole and 16-byte aligned (W % 4 == 0), coalesced scalar stores otherwise
        __syncwarp();
        if (L.d_rgb)
        {
            uint32_t const x0 = bx * kSppTileW;
            float* const dst = L.d_rgb + (static_cast<size_t>(y * W + x0) - static_cast<size_t>(L.params.buf_y0) * W) * 3;
            bool const whole = g_begin == 0 && g_end == kSppTileW / PIX;
            if (whole && x0 + kSppTileW <= W && (reinterpret_cast<uintptr_t>(dst) & 15u) == 0)
            {
                if (lane < 6) reinterpret_cast<float4*>(dst)[lane] = reinterpret_cast<float4 const*>(s_out[warp])[lane];
            }
            else
            {
                uint32_t const f = g_begin * PIX * 3u + lane;      // float index within the row segment
                if (f < g_end * PIX * 3u && x0 + f / 3u < W) dst[f] = s_out[warp][f];
            }
        }
    }
    // The item ends when its SLOWEST warp does (rows of a tile differ a lot: timing warp 0 alone missed up to
    // 180 us per item).  What it took goes to the next frame's work list; parts of a tile add up.
    if (persistent || timed_item || L.d_trace != nullptr) __syncthreads();
    if (persistent && threadIdx.x == 0) s_item = next_item;
    if (timed_item && threadIdx.x == 0 && bx < L.grid.tiles_w)
        atomicAdd(L.grid.tile_time + trace_tile, static_cast<uint32_t>(min(trace_now() - t_block, 0xffffffffull)));
    if (L.d_trace != nullptr && warp == 0)    // warp 0's busiest lane and lane sum (diagnostic)
    {
        uint32_t const db = static_cast<uint32_t>(n_box - box0), dt = static_cast<uint32_t>(n_tri - tri0);
        uint32_t const mb = __reduce_max_sync(0xffffffffu, db), mt = __reduce_max_sync(0xffffffffu, dt);
        uint32_t const sb = __reduce_add_sync(0xffffffffu, db), st = __reduce_add_sync(0xffffffffu, dt);
        trace_block(L, t_block, trace_tile, trace_cnt, (static_cast<unsigned long long>(mb) << 32) | sb,
                    (static_cast<unsigned long long>(mt) << 20) | st, n_shadow - shadow0);
    }
    if (!persistent) break;
    }
    flush_counters(L, n_primary, n_shadow, n_box, n_tri);
    if (L.d_trace != nullptr && threadIdx.x == 0) atomicMax(L.d_trace + 13, trace_now());    // latest block exit
}

// ---------------------------------------------------------------------------
// FP32 peak microbenchmark: 8 independent FFMA chains per thread.
// ---------------------------------------------------------------------------
constexpr int kPeakIters = 4096;
constexpr int kPeakChains = 8;

__global__ void __launch_bounds__(256) fp32_peak_kernel(float* out, float a, float b)
{
    float x[kPeakChains];
#pragma unroll
    for (int i = 0; i < kPeakChains; i++) x[i] = static_cast<float>(threadIdx.x + i) * 1e-3f;
#pragma unroll 1
    for (int it = 0; it < kPeakIters; it++)
    {
#pragma unroll
        for (int u = 0; u < 4; u++)
#pragma unroll
            for (int i = 0; i < kPeakChains; i++) x[i] = __fmaf_rn(x[i], a, b);
    }
    float sum = 0.f;
#pragma unroll
    for (int i = 0; i < kPeakChains; i++) sum += x[i];
    if (sum == 12345.678f) out[0] = sum; // never true; keeps the chains alive
}

} // namespace

// ---------------------------------------------------------------------------
// host launchers
// ---------------------------------------------------------------------------
// The grid serves the sample-parallel kernel with one pixel per warp pass (spp >= 32).
static int SppPixelsPerWarp(srt_render_params const& p)
{
    static int const pix_env = [] {
        char const* e = std::getenv("SRT_BVH_PIX");
        return e ? std::atoi(e) : -1;
    }();
    int pix = pix_env >= 0 ? pix_env : (p.spp >= 32 ? 1 : p.spp >= 16 ? 2 : p.spp >= 8 ? 4 : 0);
    if (pix != 0 && pix != 1 && pix != 2 && pix != 4 && pix != 8) pix = 0;
    return pix;
}

// Bit 0: the launch reads the screen grid (thread-per-pixel kernel and the one-pixel-per-warp sample-parallel
// kernel); bit 1: it reads the light grid (every BVH kernel, when shadow rays are cast).
int RenderUsesGrid(srt_render_params const& p)
{
    if (p.accel != SRT_ACCEL_BVH) return 0;
    int const pix = SppPixelsPerWarp(p);
    return ((pix == 0 || pix == 1) ? 1 : 0) | (p.shade == SRT_SHADE_SHADOW ? 2 : 0);
}

void BuildScreenGrid(GridBuild const& build, cudaStream_t stream)
{
    uint32_t const tiles = build.tiles_w * build.tiles_h;
    // tile_cnt is all zero here: zero-filled at allocation, re-zeroed by grid_scan_blocks_kernel of the last build.
    if (!build.flags_prezeroed) ML_CUDA_CHECK(cudaMemsetAsync(build.flags, 0, sizeof(GridFlags), stream));
    uint32_t const scan_blocks = (tiles + kScanCells - 1u) / kScanCells;
    uint32_t const blocks = (build.scene.n_tris_padded + 255u) / 256u;
    grid_count_kernel<<<blocks, 256, 0, stream>>>(build);
    grid_block_sums_kernel<<<scan_blocks, kScanBlock, 0, stream>>>(build);
    grid_scan_blocks_kernel<<<scan_blocks, kScanBlock, 0, stream>>>(build);
    grid_fill_kernel<<<blocks, 256, 0, stream>>>(build);
    ML_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------
// Device-side flags: the consumer half of RenderLaunch::d_signal
// ---------------------------------------------------------------------------
namespace {
__device__ __forceinline__ unsigned long long global_timer_ns()
{
    unsigned long long t;
    asm volatile("mov.u64 %0, %globaltimer;" : "=l"(t));
    return t;
}

// Thread k spins (acquire loads at system scope) until flags[k * stride] has reached `value` (wrap-safe
// signed difference); gives up after timeout_ns and counts the timeout in *timeouts (mapped host memory),
// so a dead peer can never hang the GPU.  Work enqueued behind this kernel sees everything the signalling
// kernels wrote before their release stores.
__global__ void wait_flags_kernel(uint32_t const* flags, uint32_t count, uint32_t stride, uint32_t value,
                                  unsigned long long timeout_ns, uint32_t* timeouts)
{
    uint32_t const k = blockIdx.x * blockDim.x + threadIdx.x;
    if (k >= count) return;
    uint32_t const* const f = flags + static_cast<size_t>(k) * stride;
    unsigned long long const t0 = global_timer_ns();
    uint32_t spins = 0;
    while (true)
    {
        uint32_t v;
        asm volatile("ld.acquire.sys.global.u32 %0, [%1];" : "=r"(v) : "l"(f) : "memory");
        if (static_cast<int32_t>(v - value) >= 0) break;
        if ((++spins & 63u) == 0u && global_timer_ns() - t0 > timeout_ns)
        {
            atomicAdd_system(timeouts, 1u);
            break;
        }
        __nanosleep(100);
    }
    __threadfence_system();
}

__global__ void write_flag_kernel(uint32_t* flag, uint32_t value)
{
    __threadfence_system();
    asm volatile("st.release.sys.global.u32 [%0], %1;" ::"l"(flag), "r"(value) : "memory");
}
} // namespace

void LaunchTriNormals(F4 const* e1, F4 const* e2, F4* n, uint32_t count, cudaStream_t stream)
{
    if (count == 0) return;
    tri_normals_kernel<<<(count + 255u) / 256u, 256, 0, stream>>>(reinterpret_cast<float4 const*>(e1),
                                                                 reinterpret_cast<float4 const*>(e2),
                                                                 reinterpret_cast<float4*>(n), count);
    ML_CUDA_CHECK(cudaGetLastError());
}

void LaunchSortCellsByDepth(GridBuild const& build, cudaStream_t stream)
{
    uint32_t const cells = build.tiles_w * build.tiles_h;
    sort_cells_by_depth_kernel<<<(cells + 255u) / 256u, 256, 0, stream>>>(build);
    ML_CUDA_CHECK(cudaGetLastError());
}

void LaunchWaitFlags(uint32_t const* d_flags, uint32_t count, uint32_t stride, uint32_t value,
                     uint32_t timeout_ms, uint32_t* timeouts_mapped, cudaStream_t stream)
{
    if (count == 0) return;
    uint32_t const threads = count < 256u ? count : 256u;
    wait_flags_kernel<<<(count + threads - 1u) / threads, threads, 0, stream>>>(
        d_flags, count, stride, value, static_cast<unsigned long long>(timeout_ms) * 1000000ull, timeouts_mapped);
    ML_CUDA_CHECK(cudaGetLastError());
}

void LaunchWriteFlag(uint32_t* d_flag, uint32_t value, cudaStream_t stream)
{
    write_flag_kernel<<<1, 1, 0, stream>>>(d_flag, value);
    ML_CUDA_CHECK(cudaGetLastError());
}

void LaunchRender(RenderLaunch const& launch, int sm_count, cudaStream_t stream)
{
    srt_render_params const& p = launch.params;
    uint32_t const rows = launch.n_row_blocks * 8u; // owned rows (last block may be clipped by y1)
    if (p.accel == SRT_ACCEL_BVH)
    {
        static_assert(kBvhBlockH == 8, "row blocks are 8 rows");
        dim3 grid((p.width + kBvhBlockW - 1) / kBvhBlockW, launch.n_row_blocks);
        // SRT_BVH_MINB=4|5|6: occupancy experiment (register cap 64 / 48 / 40).
        static int const minb = [] {
            char const* e = std::getenv("SRT_BVH_MINB");
            return e ? std::atoi(e) : 4;
        }();
        // Sample-parallel kernel when there are enough samples per pixel to fill the
        // lanes; SRT_BVH_PIX=1|2|4|8 pixels per warp pass (0 = pixel-per-thread kernel).
        int const pix = SppPixelsPerWarp(p);
        if (pix != 0)
        {
            dim3 sgrid((p.width + kSppTileW - 1) / kSppTileW, launch.n_row_blocks);
            // ordered launch: persistent blocks, one per resident slot (4 per SM), pulling work items
            if (pix == 1 && launch.ordered)
                sgrid = dim3(std::min<uint32_t>(sgrid.x * sgrid.y, static_cast<uint32_t>(sm_count) * 4u), 1);
            bool const st = p.collect_stats != 0;
#define SRT_SPP_LAUNCH(PIX)                                                                   \
    do {                                                                                      \
        if (st) render_bvh_spp_kernel<true, PIX><<<sgrid, 256, 0, stream>>>(launch);          \
        else render_bvh_spp_kernel<false, PIX><<<sgrid, 256, 0, stream>>>(launch);            \
    } while (0)
            static int const spp_minb = [] {
                char const* e = std::getenv("SRT_SPP_MINB");
                return e ? std::atoi(e) : 4; // 64 registers, 4 blocks per SM: with the grid paths 48 registers spill (+2.7 %, measured)
            }();
            if (pix == 1 && launch.ordered)
            {
                if (st) render_bvh_spp_kernel<true, 1, 4, true><<<sgrid, 256, 0, stream>>>(launch);
                else render_bvh_spp_kernel<false, 1, 4, true><<<sgrid, 256, 0, stream>>>(launch);
            }
            else if (pix == 1 && !st && spp_minb == 5) render_bvh_spp_kernel<false, 1, 5><<<sgrid, 256, 0, stream>>>(launch);
            else if (pix == 1 && !st && spp_minb == 6) render_bvh_spp_kernel<false, 1, 6><<<sgrid, 256, 0, stream>>>(launch);
            else if (pix == 1 && !st && spp_minb == 3) render_bvh_spp_kernel<false, 1, 3><<<sgrid, 256, 0, stream>>>(launch);
            else if (pix == 1) SRT_SPP_LAUNCH(1);
            else if (pix == 2) SRT_SPP_LAUNCH(2);
            else if (pix == 4) SRT_SPP_LAUNCH(4);
            else SRT_SPP_LAUNCH(8);
#undef SRT_SPP_LAUNCH
        }
        else if (p.collect_stats)
            render_bvh_kernel<true, 4><<<grid, kBvhThreads, 0, stream>>>(launch);
        else if (minb == 6)
            render_bvh_kernel<false, 6><<<grid, kBvhThreads, 0, stream>>>(launch);
        else if (minb == 5)
            render_bvh_kernel<false, 5><<<grid, kBvhThreads, 0, stream>>>(launch);
        else
            render_bvh_kernel<false, 4><<<grid, kBvhThreads, 0, stream>>>(launch);
    }
    else
    {
        // R rays per thread: 4 amortises the triangle LDS best; fall back to
        // fewer when the band is too small to fill the SMs twice over.
        uint64_t const pixels = static_cast<uint64_t>(rows) * p.width;
        uint64_t const fill = static_cast<uint64_t>(sm_count) * 2u * kBruteThreads;
        if (pixels >= 4 * fill)
        {
            uint32_t blocks = static_cast<uint32_t>((pixels + kBruteThreads * 4 - 1) / (kBruteThreads * 4));
            render_brute_kernel<4><<<blocks, kBruteThreads, 0, stream>>>(launch);
        }
        else if (pixels >= 2 * fill)
        {
            uint32_t blocks = static_cast<uint32_t>((pixels + kBruteThreads * 2 - 1) / (kBruteThreads * 2));
            render_brute_kernel<2><<<blocks, kBruteThreads, 0, stream>>>(launch);
        }
        else
        {
            uint32_t blocks = static_cast<uint32_t>((pixels + kBruteThreads - 1) / kBruteThreads);
            render_brute_kernel<1><<<blocks, kBruteThreads, 0, stream>>>(launch);
        }
    }
    ML_CUDA_CHECK(cudaGetLastError());
}

// ---------------------------------------------------------------------------
// srtRenderDeviceNoAlloc: the render with its per-frame scratch in STATIC device memory
// ---------------------------------------------------------------------------
// No counterpart in the reference; added for callers that cannot call
// cudaMalloc (CUDA graph capture, tight serving loops).  The scratch every render call needs -- ray counters and
// block tickets, and the screen grid's tile counts / offsets / cursors / lists / per-triangle boxes and edge
// functions -- lives in one file-scope array that every GPU loading the library holds until exit.
// Layout (regions 256-byte aligned); sizes follow SRT_NOALLOC_MAX_TRIANGLES / SRT_NOALLOC_MAX_TILES:
constexpr size_t kNaTris = SRT_NOALLOC_MAX_TRIANGLES, kNaTiles = SRT_NOALLOC_MAX_TILES;
constexpr size_t kNaAlign = 256;
constexpr size_t NaRound(size_t n) { return (n + kNaAlign - 1) / kNaAlign * kNaAlign; }
constexpr size_t kNaFrameBytes = 256;                                        // counters | tickets | GridFlags (as Scene's scratch)
constexpr size_t kNaGlobalOff = kNaFrameBytes;                               // kGridMaxGlobal u32
constexpr size_t kNaTilesOff = NaRound(kNaGlobalOff + kGridMaxGlobal * 4);   // tile_cnt | tile_off | cursor | block_sums
constexpr size_t kNaTileWords = 3 * kNaTiles + 16 + kNaTiles / kGridScanCells + 8;
constexpr size_t kNaItemsOff = NaRound(kNaTilesOff + kNaTileWords * 4);
constexpr size_t kNaItems = 16 * kNaTris;                                    // list entries; more: flags.ok = 0, BVH walk
constexpr size_t kNaBoxOff = NaRound(kNaItemsOff + kNaItems * 4);            // bbox | edges: 4 float4 per triangle slot
constexpr size_t kNaBytes = NaRound(kNaBoxOff + 4 * kNaTris * sizeof(float4));
static_assert(kNaTris % kTriTile == 0, "padded triangle count must fit");
static_assert(kNaBytes <= (size_t(8) << 20), "keep the static scratch modest");

// In host code this name is only a stand-in for the symbol: the device address comes from cudaGetSymbolAddress.
static __device__ __align__(256) unsigned char g_noalloc_scratch[kNaBytes];

// Device address of the scratch on `device` (current): resolved once per device, return code checked.
static unsigned char* NoAllocScratchAddress(int device)
{
    static std::mutex mutex;
    static unsigned char* address[64] = {};
    if (device < 0 || device >= 64) throw std::runtime_error("srtRenderDeviceNoAlloc: device ordinal out of range");
    std::lock_guard<std::mutex> lock(mutex);
    if (address[device] == nullptr)
    {
        void* ptr = nullptr;
        ML_CUDA_CHECK(cudaGetSymbolAddress(&ptr, g_noalloc_scratch));
        if (ptr == nullptr) throw std::runtime_error("srtRenderDeviceNoAlloc: cudaGetSymbolAddress returned NULL");
        address[device] = static_cast<unsigned char*>(ptr);
    }
    return address[device];
}

// `launch` / `build` arrive complete except for every pointer into scratch memory, which is set here.  Calls
// nothing that allocates: cudaGetSymbolAddress (first call per device), cudaMemsetAsync, kernel launches.
// The array keeps the previous call's contents: what must start at zero is cleared on `stream` first.  Calls
// that overlap on the GPU (other streams, other scenes, other threads) share the array and race.
void RenderNoAllocScratch(int device, RenderLaunch launch, GridBuild build, bool use_grid, int sm_count,
                          cudaStream_t stream)
{
    unsigned char* const base = NoAllocScratchAddress(device);
    unsigned long long* const frame = reinterpret_cast<unsigned long long*>(base);
    ML_CUDA_CHECK(cudaMemsetAsync(base, 0, kNaFrameBytes, stream));
    launch.d_counters = frame;                                         // 4 x u64
    launch.d_ticket = reinterpret_cast<uint32_t*>(frame) + 8;          // 2 x u32
    launch.d_signal = nullptr;
    launch.d_trace = nullptr;
    launch.grid = GridView{};
    launch.lgrid = LightGridView{};                                    // shadow rays walk the BVH: no light-grid scratch
    if (use_grid)
    {
        uint32_t* const t = reinterpret_cast<uint32_t*>(base + kNaTilesOff);
        uint32_t const tiles = build.tiles_w * build.tiles_h;
        ML_CUDA_CHECK(cudaMemsetAsync(t, 0, static_cast<size_t>(tiles) * sizeof(uint32_t), stream));   // tile counts
        build.cap = static_cast<uint32_t>(kNaItems);
        build.tile_cnt = t; build.tile_off = t + kNaTiles; build.cursor = t + 2 * kNaTiles + 1;
        build.block_sums = t + 3 * kNaTiles + 4;
        build.items = reinterpret_cast<uint32_t*>(base + kNaItemsOff);
        build.bbox = reinterpret_cast<F4*>(base + kNaBoxOff);
        build.edges = build.bbox + build.scene.n_tris_padded;
        build.flags = reinterpret_cast<GridFlags*>(reinterpret_cast<uint32_t*>(frame) + 12);
        build.flags_prezeroed = 1;
        build.global = reinterpret_cast<uint32_t*>(base + kNaGlobalOff);
        BuildScreenGrid(build, stream);
        launch.grid.tile_off = build.tile_off; launch.grid.items = build.items; launch.grid.bbox = build.bbox;
        launch.grid.edges = build.edges; launch.grid.global = build.global; launch.grid.flags = build.flags;
        launch.grid.tiles_w = build.tiles_w; launch.grid.tiles_h = build.tiles_h;
        launch.grid.enabled = 1;
    }
    if (launch.n_row_blocks > 0) LaunchRender(launch, sm_count, stream);
}

double MeasureFp32Peak(DeviceState const& device, int repeats)
{
    device.MakeCurrent();
    float* d_out = nullptr;
    ML_CUDA_CHECK(cudaMalloc(&d_out, sizeof(float)));
    int const blocks = device.prop.multiProcessorCount * 16;
    cudaStream_t s = device.stream;
    fp32_peak_kernel<<<blocks, 256, 0, s>>>(d_out, 0.999f, 1e-3f); // warm-up
    double best_ms = 1e30;
    for (int r = 0; r < (repeats > 0 ? repeats : 5); r++)
    {
        ML_CUDA_CHECK(cudaEventRecord(device.ev_start, s));
        fp32_peak_kernel<<<blocks, 256, 0, s>>>(d_out, 0.999f, 1e-3f);
        ML_CUDA_CHECK(cudaEventRecord(device.ev_stop, s));
        ML_CUDA_CHECK(cudaStreamSynchronize(s));
        float ms = 0.f;
        ML_CUDA_CHECK(cudaEventElapsedTime(&ms, device.ev_start, device.ev_stop));
        if (ms < best_ms) best_ms = ms;
    }
    cudaFree(d_out);
    double flops = 2.0 * kPeakIters * 4.0 * kPeakChains * 256.0 * blocks;
    return flops / (best_ms * 1e-3) / 1e12;
}

} // namespace ML
