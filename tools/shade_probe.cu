// shade_probe.cu -- what the FP64 island of a hit's shading costs on sm_100a, instruction by instruction and function by
// function (the deviates of RandomDiretion, the draws, the sky's pow).  Includes the product's own csrc/rtc_math.cuh.
//
//   nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iraytracingc_b200/csrc -Iinclude -o /tmp/shade_probe tools/shade_probe.cu
//
// Every number is in CLOCKS OF ONE SM SUB-PARTITION (clock64 inside the kernel, so independent of the clock frequency):
//   "issue"   = clocks per warp-level call with 8 warps resident per sub-partition (throughput: what a call costs when
//               other warps fill the gaps -- the pipe or issue-slot occupancy);
//   "latency" = clocks per call with ONE warp per sub-partition (what a warp waits when nothing else can run).
#include <cstdio>
#include <cstdlib>
#include <cuda_runtime.h>

#include "rtc_math.cuh"

using namespace rtc;

// ---- candidate forms of RandomValue()'s (float)(r / 4294967295.0), all verified exhaustively on the CPU ------------
__device__ __forceinline__ float unit_magic(uint32_t r) // the conversion from the 2^52 trick: DADD instead of I2F.F64
{
    const double q = __hiloint2double(0x43300000, (int)r) - 4503599627370496.0;
    return (float)(__fma_rn(q, 2.3283064365386963e-10, q) * 2.3283064365386963e-10);
}
__device__ __forceinline__ float unit_f32(uint32_t r) // FP32 only: two exact I2FP, two FFMA (ties pushed up), the 4 double-rounding cases
{
    const float fh = __uint2float_rn(r >> 8), fl = __uint2float_rn(r & 255u);
    const float c = r >= 0xFFFFFC00u ? __fadd_rn(fl, 1.0f) : __fmaf_rn(fh, 0x1p-30f, fl);
    return __fmul_rn(__fmaf_rn(fh, 256.0f, c), 0x1p-32f);
}

template <int MODE> __device__ __forceinline__ float body(uint32_t &st, float acc)
{
    if (MODE == 0) { const float a = random_value(st), b = random_value(st); return acc + normal_from_draws_exact(a, b); }
    if (MODE == 1) { const float a = random_value(st), b = random_value(st); return acc + normal_from_draws(a, b); }
    if (MODE == 2) { const float a = random_value(st), b = random_value(st); return acc + a * b; }
    if (MODE == 3) { const float a = unit_magic(rng_next(st)), b = unit_magic(rng_next(st)); return acc + a * b; }
    if (MODE == 4) { const float a = unit_f32(rng_next(st)), b = unit_f32(rng_next(st)); return acc + a * b; }
    if (MODE == 5) { const float a = random_value(st); return acc + xpowf(a, 0.35f); }
    if (MODE == 8) { const float a = random_value(st); return acc + (float)exp2(0.35 * log2((double)a)); } // round 1's form
    if (MODE == 6) { const float4 d = random_direction(st); st = __float_as_uint(d.w); return acc + d.x; }
    if (MODE == 7) { const uint32_t a = rng_next(st), b = rng_next(st); return acc + __uint_as_float((a ^ b) >> 9 | 0x3f800000u); } // the generator alone
    return acc;
}

template <int MODE> __global__ void __launch_bounds__(512, 2) fn_probe(float *out, long long *clk, int iters)
{
    uint32_t st = (blockIdx.x * blockDim.x + threadIdx.x) * 2654435761u + 12345u;
    float acc = 0.f;
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; ++i)
        acc = body<MODE>(st, acc);
    const long long t1 = clock64();
    out[blockIdx.x * blockDim.x + threadIdx.x] = acc;
    if (threadIdx.x == 0)
        clk[blockIdx.x] = t1 - t0;
}

// ---- single instructions: 8 independent chains per thread ----------------------------------------------------------
template <int MODE> __global__ void __launch_bounds__(512, 2) op_probe(float *out, long long *clk, int iters, double da, float fa)
{
    double d[8];
    float f[8];
    uint32_t u[8];
#pragma unroll
    for (int k = 0; k < 8; ++k) {
        d[k] = 1.0 + 1e-3 * (threadIdx.x + k);
        f[k] = 1.0f + 1e-3f * (threadIdx.x + k);
        u[k] = threadIdx.x * 977u + k;
    }
    const long long t0 = clock64();
#pragma unroll 1
    for (int i = 0; i < iters; ++i) {
#pragma unroll
        for (int k = 0; k < 8; ++k) {
            if (MODE == 0) d[k] = fma(d[k], da, da);                                         // DFMA
            if (MODE == 1) d[k] = __dadd_rn(d[k], da);                                       // DADD
            if (MODE == 2) d[k] = __dmul_rn(d[k], da);                                       // DMUL
            if (MODE == 3) { f[k] = (float)d[k]; d[k] = __hiloint2double(__double2hiint(d[k]), __float_as_int(f[k])); } // F2F.F32.F64
            if (MODE == 4) { d[k] = (double)f[k]; f[k] = __int_as_float(__double2loint(d[k]) | 0x3f800000); }            // F2F.F64.F32
            if (MODE == 5) { d[k] = (double)u[k]; u[k] ^= (uint32_t)__double2hiint(d[k]); }  // I2F.F64.U32
            if (MODE == 6) { f[k] = __uint2float_rn(u[k] & 0xffffffu); u[k] += __float_as_uint(f[k]); } // I2FP.F32.U32 (+ LOP + IADD)
            if (MODE == 7) asm volatile("rcp.approx.ftz.f32 %0, %0;" : "+f"(f[k]));         // MUFU.RCP
            if (MODE == 8) d[k] = rint(d[k] * da);                                           // FRND.F64 (+ DMUL)
            if (MODE == 9) u[k] = (u[k] + 0x10000000u) & (u[(k + 1) & 7] | 0xE0000000u);     // IADD + LOP3 (the integer narrowing)
            if (MODE == 10) f[k] = __fmaf_rn(f[k], fa, fa);                                  // FFMA
        }
    }
    const long long t1 = clock64();
    float s = 0.f;
#pragma unroll
    for (int k = 0; k < 8; ++k)
        s += (float)d[k] + f[k] + (float)u[k];
    out[blockIdx.x * blockDim.x + threadIdx.x] = s;
    if (threadIdx.x == 0)
        clk[blockIdx.x] = t1 - t0;
}

static float *g_out;
static long long *g_clk;
static int g_sms;

template <typename K, typename... A> static double run(K kern, int threads, int iters, A... extra)
{
    // 512 threads x 2 blocks per SM = 8 warps per sub-partition, resident together (64 registers, no shared memory);
    // 128 threads x 1 block = one warp per sub-partition
    const int blocks_per_sm = threads == 128 ? 1 : 2;
    const int grid = g_sms * blocks_per_sm;
    kern<<<grid, threads>>>(g_out, g_clk, iters, extra...); // warm-up
    kern<<<grid, threads>>>(g_out, g_clk, iters, extra...);
    if (cudaDeviceSynchronize() != cudaSuccess) {
        printf("launch failed: %s\n", cudaGetErrorString(cudaGetLastError()));
        exit(1);
    }
    static long long h[4096];
    cudaMemcpy(h, g_clk, sizeof(long long) * grid, cudaMemcpyDeviceToHost);
    double sum = 0;
    for (int i = 0; i < grid; ++i)
        sum += (double)h[i];
    const double clocks = sum / grid;                      // per block, all blocks run concurrently
    const int warps_per_smsp = threads * blocks_per_sm / 128;
    return clocks / ((double)iters * warps_per_smsp);      // sub-partition clocks per warp-level iteration
}

#define FN(MODE, name)                                                                                              \
    {                                                                                                               \
        const double thr = run(fn_probe<MODE>, 512, 2000), lat = run(fn_probe<MODE>, 128, 2000);                    \
        printf("%-58s issue %7.1f   latency %8.1f\n", name, thr, lat);                                              \
    }
#define OP(MODE, name)                                                                                              \
    {                                                                                                               \
        const double thr = run(op_probe<MODE>, 512, 1000, 1.0000001, 1.0000001f) / 8.0;                             \
        printf("%-58s %6.2f clocks per warp instruction\n", name, thr);                                             \
    }

int main()
{
    cudaDeviceProp p;
    cudaGetDeviceProperties(&p, 0);
    g_sms = p.multiProcessorCount;
    cudaMalloc(&g_out, sizeof(float) * 4096 * 512);
    cudaMalloc(&g_clk, sizeof(long long) * 4096);
    printf("%s, %d SMs; clocks of one SM sub-partition per warp-level call\n", p.name, g_sms);
    FN(7, "generator alone (2 x rng_next)");
    FN(2, "2 draws, product form (I2F.F64, DFMA, DMUL, F2F)");
    FN(3, "2 draws, 2^52 trick instead of I2F.F64");
    FN(4, "2 draws, FP32 only (2 I2FP, 2 FFMA, FMUL, select)");
    FN(0, "2 draws + normal deviate, exact sequence (libm)");
    FN(1, "2 draws + normal deviate, fast path + fallback");
    FN(6, "random_direction (3 deviates + normalize), product build");
    FN(5, "1 draw + xpowf(u, .35): glibc's algorithm restated");
    FN(8, "1 draw + (float)exp2(.35 * log2(u)), CUDA's double libm");
    OP(10, "FFMA");
    OP(0, "DFMA");
    OP(1, "DADD");
    OP(2, "DMUL");
    OP(3, "F2F.F32.F64 (+ int op)");
    OP(4, "F2F.F64.F32 (+ int op)");
    OP(5, "I2F.F64.U32 (+ int op)");
    OP(6, "I2FP.F32.U32 (+ 2 int ops)");
    OP(7, "MUFU.RCP");
    OP(8, "FRND.F64 + DMUL");
    OP(9, "IADD + LOP3");
    return 0;
}
