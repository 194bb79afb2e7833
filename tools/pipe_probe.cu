// pipe_probe.cu -- issue-rate micro-benchmarks for the instruction mixes of the render kernels (sm_100a).
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -o tools/pipe_probe.bin tools/pipe_probe.cu
// Prints warp-instructions per clock per SM for each mix (4.0 = one instruction per SMSP per clock).
#include <cstdio>
#include <cuda_fp16.h>
#include <cuda_runtime.h>

#define ITERS 2048
#define NCH 8

template <int MODE>
__global__ void __launch_bounds__(256) probe(float *out, float a, float b, int n)
{
    float r[NCH];
    unsigned u[NCH];
    __half2 h[NCH], hs = __floats2half2_rn(a, b);
    unsigned long long p2[NCH], bb;
    asm("mov.b64 %0, {%1, %2};" : "=l"(bb) : "f"(a), "f"(b + threadIdx.x));
#pragma unroll
    for (int k = 0; k < NCH; ++k) {
        r[k] = threadIdx.x + k;
        u[k] = threadIdx.x * 977u + k;
        h[k] = __floats2half2_rn(0.001f * threadIdx.x, 0.5f * k);
        asm("mov.b64 %0, {%1, %2};" : "=l"(p2[k]) : "f"(0.5f + 1e-3f * k), "f"(0.25f + 1e-3f * threadIdx.x));
    }
    for (int i = 0; i < n; ++i) {
#pragma unroll
        for (int k = 0; k < NCH; ++k) {
            if (MODE == 0) r[k] = fmaf(r[k], a, b);                       // FFMA reg, const, const
            if (MODE == 1) r[k] = __fadd_rn(r[k], r[(k + 1) % NCH]);      // FADD reg-reg
            if (MODE == 2) h[k] = __hadd2(h[k], hs);                      // HADD2
            if (MODE == 3) u[k] = (u[k] | u[(k + 1) % NCH]) & u[(k + 2) % NCH]; // LOP3
            if (MODE == 4) {                                              // 4 FADD + 2 LOP3 (current filter, per pixel)
                r[k] = __fadd_rn(r[k], a);
                if ((k & 1) == 0) u[k] = (u[k] | __float_as_uint(r[k])) & u[(k + 1) % NCH];
            }
            if (MODE == 5) {                                              // 2 HADD2 + 1 LOP3 (packed filter, per pixel)
                h[k] = __hadd2(h[k], hs);
                if ((k & 1) == 0) {
                    unsigned x = *reinterpret_cast<unsigned *>(&h[k]), y = *reinterpret_cast<unsigned *>(&h[k + 1]);
                    u[k] &= (x | y);
                }
            }
            if (MODE == 6) r[k] = __fmul_rn(r[k], r[(k + 1) % NCH]);      // FMUL reg-reg
            if (MODE == 7) r[k] = fmaf(r[k], r[(k + 1) % NCH], r[(k + 2) % NCH]); // FFMA 3-reg
            if (MODE == 8) {                                              // FMUL + FADD alternating (exact dot products)
                r[k] = (k & 1) ? __fmul_rn(r[k], r[(k + 1) % NCH]) : __fadd_rn(r[k], r[(k + 1) % NCH]);
            }
            if (MODE == 9) u[k] = max(u[k], u[(k + 1) % NCH]);            // IMNMX / VIMNMX (alu)
            if (MODE == 10) r[k] = fmaxf(r[k], r[(k + 1) % NCH]);         // FMNMX (alu)
            if (MODE == 11) asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2[k]) : "l"(p2[(k + 1) % NCH]), "l"(p2[(k + 2) % NCH])); // FFMA2 3-reg
            if (MODE == 17) asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p2[k]) : "l"(p2[(k + 1) % NCH]), "l"(bb)); // FFMA2 acc, reg64, SAME reg64 (reuse)
            if (MODE == 18) { unsigned long long bk; asm("mov.b64 %0, {%1, %1};" : "=l"(bk) : "f"(r[k])); // FFMA2 acc, reg64, per-k .F32 broadcast
                asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p2[k]) : "l"(p2[(k + 1) % NCH]), "l"(bk)); }
            if (MODE == 19) { unsigned long long bk; asm("mov.b64 %0, {%1, %1};" : "=l"(bk) : "f"(r[0])); // FFMA2 acc, reg64, SAME .F32 broadcast
                asm volatile("fma.rn.f32x2 %0, %1, %2, %0;" : "+l"(p2[k]) : "l"(p2[(k + 1) % NCH]), "l"(bk)); }
            if (MODE == 20) r[k] = fmaf(r[(k + 3) % NCH], a, r[k]);       // FFMA acc, reg, const
            if (MODE == 21) r[k] = fmaf(r[(k + 3) % NCH], r[(k + 5) % NCH], r[k]); // FFMA acc, reg, reg (other spacing)
            if (MODE == 12) {                                             // FFMA2 + FMNMX alternating (fma pipe + alu pipe)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2[k]) : "l"(p2[(k + 1) % NCH]), "l"(p2[(k + 2) % NCH]));
                r[k] = fmaxf(r[k], r[(k + 1) % NCH]);
            }
            if (MODE == 13) {                                             // FFMA + FMNMX alternating
                r[k] = fmaf(r[k], r[(k + 1) % NCH], r[(k + 2) % NCH]);
                u[k] = max(u[k], u[(k + 1) % NCH]);
            }
            if (MODE == 14) {                                             // FFMA2 + FFMA alternating (same pipe?)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2[k]) : "l"(p2[(k + 1) % NCH]), "l"(p2[(k + 2) % NCH]));
                r[k] = fmaf(r[k], r[(k + 1) % NCH], r[(k + 2) % NCH]);
            }
            if (MODE == 15) {                                             // 2 FFMA2 + 1 FSETP-like (packed sweep's mix)
                asm volatile("fma.rn.f32x2 %0, %0, %1, %2;" : "+l"(p2[k]) : "l"(p2[(k + 1) % NCH]), "l"(p2[(k + 2) % NCH]));
                if ((k & 1) == 0) asm volatile("{.reg .pred q; setp.geu.f32 q, %1, %2; @q or.b32 %0, %0, 1;}" : "+r"(u[k]) : "f"(r[k]), "f"(r[(k + 1) % NCH]));
            }
            if (MODE == 16) {                                             // 2 FFMA + 1 FSETP-like (scalar sweep's mix)
                r[k] = fmaf(r[k], a, b);
                if ((k & 1) == 0) asm volatile("{.reg .pred q; setp.geu.f32 q, %1, %2; @q or.b32 %0, %0, 1;}" : "+r"(u[k]) : "f"(r[k]), "f"(r[(k + 1) % NCH]));
            }
        }
    }
    float s = 0;
#pragma unroll
    for (int k = 0; k < NCH; ++k)
        s += r[k] + (float)u[k] + __low2float(h[k]) + __high2float(h[k]) + (float)(p2[k] >> 13);
    if (s == 123.456f)
        out[blockIdx.x * blockDim.x + threadIdx.x] = s;
}

template <int MODE>
void run(const char *name, double inst_per_iter_per_thread)
{
    int dev = 0, sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, dev);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, dev);
    float *buf;
    cudaMalloc(&buf, 4 << 20);
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    int blocks = sms * 8;
    float best = 1e30f;
    for (int rep = 0; rep < 5; ++rep) {
        cudaEventRecord(e0);
        probe<MODE><<<blocks, 256>>>(buf, 1.000001f, 0.5f, ITERS);
        cudaEventRecord(e1);
        cudaEventSynchronize(e1);
        float ms;
        cudaEventElapsedTime(&ms, e0, e1);
        if (rep && ms < best)
            best = ms;
    }
    double warp_inst = inst_per_iter_per_thread * ITERS * (256.0 / 32) * blocks;
    double clocks = best * 1e-3 * khz * 1e3;
    printf("%-44s %8.3f ms  %6.2f warp-inst/clk/SM (at %d MHz nominal)\n", name, best, warp_inst / clocks / sms,
           khz / 1000);
    cudaFree(buf);
}

int main()
{
    run<0>("FFMA r,c,c", NCH);
    run<7>("FFMA r,r,r", NCH);
    run<1>("FADD r,r", NCH);
    run<6>("FMUL r,r", NCH);
    run<8>("FMUL/FADD alternating", NCH);
    run<2>("HADD2", NCH);
    run<3>("LOP3", NCH);
    run<9>("IMNMX", NCH);
    run<10>("FMNMX", NCH);
    run<11>("FFMA2 r,r,r (warp-inst; x2 FMA each)", NCH);
    run<17>("FFMA2 acc, r64, same r64", NCH);
    run<18>("FFMA2 acc, r64, per-instr .F32 broadcast", NCH);
    run<19>("FFMA2 acc, r64, same .F32 broadcast", NCH);
    run<20>("FFMA acc, r, c", NCH);
    run<21>("FFMA acc, r, r", NCH);
    run<12>("FFMA2 + FMNMX alternating", 2 * NCH);
    run<13>("FFMA + IMNMX alternating", 2 * NCH);
    run<14>("FFMA2 + FFMA alternating", 2 * NCH);
    run<15>("2 FFMA2 + (FSETP + @p LOP)", NCH + NCH);
    run<16>("2 FFMA + (FSETP + @p LOP)", NCH + NCH);
    run<4>("4 FADD + 2 LOP3 (x2)", NCH + NCH / 2);
    run<5>("2 HADD2 + 1 LOP3 (x4)", NCH + NCH / 2);
    return 0;
}
