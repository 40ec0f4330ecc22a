// ubench_fp32.cu -- FP32 issue/pipe microbenchmarks for B200 (sm_100a).
// Answers, on the box: (1) scalar FFMA peak; (2) packed fma.rn.f32x2 (SASS FFMA2)
// peak in FLOP/s and in issue slots; (3) whether ALU-pipe instructions (LOP3)
// co-issue next to FFMA / FFMA2 without costing FMA throughput.
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o ubench_fp32 ubench_fp32.cu
#include <cstdio>
#include <cuda_runtime.h>

constexpr int kIters = 2048;

__device__ __forceinline__ unsigned long long pack(float a, float b)
{
    unsigned long long r;
    asm("mov.b64 %0, {%1, %2};" : "=l"(r) : "f"(a), "f"(b));
    return r;
}
__device__ __forceinline__ unsigned long long fma2(unsigned long long a, unsigned long long b, unsigned long long c)
{
    unsigned long long d;
    asm("fma.rn.f32x2 %0, %1, %2, %3;" : "=l"(d) : "l"(a), "l"(b), "l"(c));
    return d;
}

// mode 0: 16 scalar FFMA chains; mode 1: + one LOP3 per FFMA
template <int ALU>
__global__ void __launch_bounds__(256) k_scalar(float* out, float a, float b, unsigned m)
{
    float x[16];
    unsigned u[16];
#pragma unroll
    for (int i = 0; i < 16; i++) { x[i] = threadIdx.x * 1e-3f + i; u[i] = threadIdx.x + i; }
#pragma unroll 1
    for (int it = 0; it < kIters; it++)
    {
#pragma unroll
        for (int i = 0; i < 16; i++)
        {
            x[i] = __fmaf_rn(x[i], a, b);
            if (ALU) u[i] = (u[i] ^ m) + (u[i] >> 3 & m);   // LOP3/SHF/IADD on the ALU pipe
        }
    }
    float s = 0.f; unsigned v = 0;
#pragma unroll
    for (int i = 0; i < 16; i++) { s += x[i]; v ^= u[i]; }
    if (s == 12345.678f || v == 0xdeadbeefu) out[0] = s + v;
}

template <int ALU>
__global__ void __launch_bounds__(256) k_packed(float* out, float a, float b, unsigned m)
{
    unsigned long long x[8];
    unsigned u[16];
    unsigned long long pa = pack(a, a), pb = pack(b, b);
#pragma unroll
    for (int i = 0; i < 8; i++) x[i] = pack(threadIdx.x * 1e-3f + i, threadIdx.x * 2e-3f + i);
#pragma unroll
    for (int i = 0; i < 16; i++) u[i] = threadIdx.x + i;
#pragma unroll 1
    for (int it = 0; it < kIters; it++)
    {
#pragma unroll
        for (int i = 0; i < 8; i++)
        {
            x[i] = fma2(x[i], pa, pb);
            if (ALU)
            {
                u[2 * i] = (u[2 * i] ^ m) + (u[2 * i] >> 3 & m);
                u[2 * i + 1] = (u[2 * i + 1] ^ m) + (u[2 * i + 1] >> 3 & m);
            }
        }
    }
    unsigned long long s = 0; unsigned v = 0;
#pragma unroll
    for (int i = 0; i < 8; i++) s ^= x[i];
#pragma unroll
    for (int i = 0; i < 16; i++) v ^= u[i];
    if (s == 0x123456789abcdefull || v == 0xdeadbeefu) out[0] = (float)s + v;
}

template <class K>
double time_kernel(K kernel, int blocks, float* d_out)
{
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0); cudaEventCreate(&e1);
    kernel<<<blocks, 256>>>(d_out, 0.999f, 1e-3f, 0x5555u);
    double best = 1e30;
    for (int r = 0; r < 5; r++)
    {
        cudaEventRecord(e0);
        kernel<<<blocks, 256>>>(d_out, 0.999f, 1e-3f, 0x5555u);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms; cudaEventElapsedTime(&ms, e0, e1);
        if (ms < best) best = ms;
    }
    return best;
}

int main()
{
    cudaDeviceProp prop; cudaGetDeviceProperties(&prop, 0);
    int blocks = prop.multiProcessorCount * 8;
    float* d_out; cudaMalloc(&d_out, 4);
    double fmas = 16.0 * kIters * 256.0 * blocks;   // fp32 FMAs per launch in every variant
    struct { const char* name; double ms; } r[4] = {
        {"scalar FFMA", time_kernel(k_scalar<0>, blocks, d_out)},
        {"scalar FFMA + ALU ops", time_kernel(k_scalar<1>, blocks, d_out)},
        {"packed fma.rn.f32x2", time_kernel(k_packed<0>, blocks, d_out)},
        {"packed fma.rn.f32x2 + ALU ops", time_kernel(k_packed<1>, blocks, d_out)},
    };
    printf("%s, %d SMs\n", prop.name, prop.multiProcessorCount);
    for (auto& x : r)
        printf("%-32s %8.3f ms  %7.2f TFLOP/s (fp32 FMA = 2)\n", x.name, x.ms, 2.0 * fmas / (x.ms * 1e-3) / 1e12);
    printf("cudaGetLastError: %s\n", cudaGetErrorString(cudaGetLastError()));
    return 0;
}
