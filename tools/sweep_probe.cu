// sweep_probe.cu -- the path kernel's branch-free first phase (rtc_paths.cuh) in isolation, to compare ways of
// feeding the per-triangle quick records to the FMA pipe on sm_100a:
//   0  scalar FFMA chains, records by warp-uniform LDS.128                      (MODE 1 of the kernel)
//   1  packed FFMA2 (two triangles per instruction), pair records by LDS.128    (MODE 3)
//   2  packed FFMA2, pair records from __constant__ memory (LDCU -> uniform-register operand)
//   3  scalar FFMA, records from __constant__ memory (uniform-register operand)
// Every lane sweeps its own ray over the same N triangles REPS times; prints pairs/s and issue-slots-equivalent.
// Build: nvcc -gencode arch=compute_100a,code=sm_100a -O3 -std=c++17 -Iraytracingc_b200/csrc -Iinclude \
//             -o tools/sweep_probe.bin tools/sweep_probe.cu
#include <cstdio>
#include <cstdlib>
#include <vector>

#include "rtc_paths.cuh"

using namespace rtc;

constexpr int CQ_F4 = 3072; // 48 KB of the 64 KB constant space
__constant__ float4 cq[CQ_F4];

template <int V>
__global__ void __launch_bounds__(128, 6) probe(const float4 *__restrict__ qrec, int n, int reps, unsigned *out)
{
    extern __shared__ float4 sQ[];
    const int nq = (int)path_q_f4((V == 0 || V == 3) ? 1 : (V == 6 ? 5 : 3), n);
    if (V < 2 || V >= 4) {
        for (int i = threadIdx.x; i < nq; i += blockDim.x)
            sQ[i] = qrec[i];
        __syncthreads();
    }
    const unsigned g = blockIdx.x * blockDim.x + threadIdx.x;
    // a ray inside the bounds the records were built for
    f3 o = mk3(0.37f * ((g % 11) - 5.f), 0.21f * ((g % 7) - 3.f), 0.4f * ((g % 5) - 2.f));
    f3 d = mk3(0.6f - 0.001f * (g % 97), -0.3f + 0.002f * (g % 53), 0.74f - 0.003f * (g % 31));
    unsigned acc = 0;
    for (int r = 0; r < reps; ++r) {
        const QuickRay qr = quick_ray(o, d);
        const QuickRay2 qr2 = quick_ray2(qr);
        for (int c0 = 0; c0 < n && V < 5; c0 += 32) {
            const int cn = min(32, n - c0);
            unsigned m = 0;
            if (V == 0 || V == 3) {
                int j = 0;
#pragma unroll 1
                for (; j + 4 <= cn; j += 4) {
                    unsigned mg = 0;
                    if (V == 0)
                        quick2_group<0, 4>(qr, sQ + 4 * (c0 + j), mg);
                    else {
                        const float4 *R = cq + 4 * (c0 + j);
                        tri_quick2_bit<1u>(qr, R[0], R[1], R[2], R[3], mg);
                        tri_quick2_bit<2u>(qr, R[4], R[5], R[6], R[7], mg);
                        tri_quick2_bit<4u>(qr, R[8], R[9], R[10], R[11], mg);
                        tri_quick2_bit<8u>(qr, R[12], R[13], R[14], R[15], mg);
                    }
                    m |= mg << j;
                }
            } else {
                const int np = (cn + 1) >> 1;
                int j = 0;
#pragma unroll 1
                for (; j + 2 <= np; j += 2) {
                    unsigned mg = 0;
                    if (V == 1) {
                        quick2_pairs<0, 2>(qr2, sQ + QPAIR_F4 * ((c0 >> 1) + j), mg);
                    } else if (V == 4) {
                        QuickVals v[4];
                        quick_vals2x2(qr2, sQ + QPAIR_F4 * ((c0 >> 1) + j), v);
                        quick_keep_bit<1u>(v[0], mg);
                        quick_keep_bit<2u>(v[1], mg);
                        quick_keep_bit<4u>(v[2], mg);
                        quick_keep_bit<8u>(v[3], mg);
                    } else {
                        QuickVals v0, v1;
                        quick_vals2(qr2, cq + QPAIR_F4 * ((c0 >> 1) + j), v0, v1);
                        quick_keep_bit<1u>(v0, mg);
                        quick_keep_bit<2u>(v1, mg);
                        quick_vals2(qr2, cq + QPAIR_F4 * ((c0 >> 1) + j + 1), v0, v1);
                        quick_keep_bit<4u>(v0, mg);
                        quick_keep_bit<8u>(v1, mg);
                    }
                    m |= mg << (2 * j);
                }
            }
            acc += __popc(m);
        }
        if (V == 6) { // lane = triangle, rays two at a time from a per-warp scratch (MODE 5 of the kernel, first phase only)
            const int lane = threadIdx.x & 31;
            float4 *scr = sQ + QTRI_F4 * n + PATH_SCRATCH_F4 * (threadIdx.x >> 5);
            float *sf = reinterpret_cast<float *>(scr) + 12 * (lane >> 1) + (lane & 1);
            sf[0] = qr.d.x, sf[2] = qr.d.y, sf[4] = qr.d.z, sf[6] = qr.m.x, sf[8] = qr.m.y, sf[10] = qr.m.z;
            uint2 *msk = reinterpret_cast<uint2 *>(scr + 16 * 3);
            __syncwarp();
            const f32x2 kk = pk2(1.0000019073486328f, 1.0000019073486328f);
            for (int c0 = 0; c0 < n; c0 += 32) {
                const int cn = min(32, n - c0);
                float4 r0 = make_float4(0.f, 0.f, 0.f, 0.f), r1 = r0, r2 = r0, r3 = make_float4(0.f, 0.f, 0.f, __int_as_float(0x7f800000));
                if (lane < cn) {
                    const float4 *R = sQ + QTRI_F4 * (c0 + lane);
                    r0 = R[0], r1 = R[1], r2 = R[2], r3 = R[3];
                }
                QuickVals v0, v1;
                v0.thr = v1.thr = r3.w;
                v0.M = v1.M = QUICK_EPS - r3.w;
                const f32x2 M2 = pk2(v0.M, v0.M);
                const f32x2 c0x = pk2(r0.x, r0.x), c0y = pk2(r0.y, r0.y), c0z = pk2(r0.z, r0.z), c0w = pk2(r0.w, r0.w);
                const f32x2 c1x = pk2(r1.x, r1.x), c1y = pk2(r1.y, r1.y), c1z = pk2(r1.z, r1.z), c1w = pk2(r1.w, r1.w);
                const f32x2 c2x = pk2(r2.x, r2.x), c2y = pk2(r2.y, r2.y), c2z = pk2(r2.z, r2.z), c2w = pk2(r2.w, r2.w);
                const f32x2 c3x = pk2(r3.x, r3.x), c3y = pk2(r3.y, r3.y), c3z = pk2(r3.z, r3.z);
                auto slot = [&](int k) {
                    const ulonglong2 *Q = reinterpret_cast<const ulonglong2 *>(scr + 3 * k);
                    const ulonglong2 qa = Q[0], qb = Q[1], qc = Q[2];
                    const f32x2 U = fma2(c1y, qb.x, fma2(c1x, qa.y, fma2(c0w, qa.x, fma2(c0z, qc.y, fma2(c0y, qc.x, mul2(c0x, qb.y))))));
                    const f32x2 Vv = fma2(c2w, qb.x, fma2(c2z, qa.y, fma2(c2y, qa.x, fma2(c2x, qc.y, fma2(c1w, qc.x, mul2(c1z, qb.y))))));
                    const f32x2 det = fma2(c3z, qb.x, fma2(c3y, qa.y, mul2(c3x, qa.x)));
                    const f32x2 sum = add2(U, Vv);
                    const f32x2 hi = fma2(det, kk, M2);
                    unpk2(U, v0.U, v1.U);
                    unpk2(Vv, v0.V, v1.V);
                    unpk2(det, v0.det, v1.det);
                    unpk2(sum, v0.sum, v1.sum);
                    unpk2(hi, v0.hi, v1.hi);
                    const unsigned b0 = quick_keep_ballot(v0), b1 = quick_keep_ballot(v1);
                    if (lane == 0)
                        msk[k] = make_uint2(b0, b1);
                };
#pragma unroll 1
                for (int k = 0; k < 16; k += 2) {
                    slot(k);
                    slot(k + 1);
                }
                __syncwarp();
                acc += __popc(reinterpret_cast<const unsigned *>(msk)[lane]);
                __syncwarp();
            }
        }
        if (V == 5) { // two rays per lane: every pair record is loaded once for both
            f3 o2 = mk3(o.y, o.z, o.x), d2 = mk3(d.y, d.x, d.z);
            const QuickRay2 qb = quick_ray2(quick_ray(o2, d2));
            for (int c0 = 0; c0 < n; c0 += 32) {
                const int np = (min(32, n - c0) + 1) >> 1;
                unsigned ma = 0, mb = 0;
#pragma unroll 1
                for (int j = 0; j < np; ++j) {
                    const float4 *P = sQ + QPAIR_F4 * ((c0 >> 1) + j);
                    QuickVals v0, v1, w0, w1;
                    quick_vals2(qr2, P, v0, v1);
                    quick_vals2(qb, P, w0, w1);
                    unsigned mg = 0, mh = 0;
                    quick_keep_bit<1u>(v0, mg);
                    quick_keep_bit<2u>(v1, mg);
                    quick_keep_bit<1u>(w0, mh);
                    quick_keep_bit<2u>(w1, mh);
                    ma |= mg << (2 * j);
                    mb |= mh << (2 * j);
                }
                acc += __popc(ma) + __popc(mb);
            }
        }
        o.x += 1e-4f; // the next "ray": keeps the sweep from being hoisted out of the loop
        d.y += 1e-5f;
    }
    out[g] = acc;
}

int main(int argc, char **argv)
{
    const int n = argc > 1 ? atoi(argv[1]) : 120;
    const int reps = argc > 2 ? atoi(argv[2]) : 400;
    int sms = 0, khz = 0;
    cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    // random triangles of the bundled scenes' scale
    std::vector<float4> geo(3 * n);
    srand(7);
    auto rnd = [](float a, float b) { return a + (b - a) * (rand() / (float)RAND_MAX); };
    for (int i = 0; i < n; ++i) {
        float A[3] = {rnd(-2, 2), rnd(-2, 2), rnd(-2, 2)}, e1[3] = {rnd(-1, 1), rnd(-1, 1), rnd(-1, 1)},
              e2[3] = {rnd(-1, 1), rnd(-1, 1), rnd(-1, 1)};
        float nx = e1[1] * e2[2] - e1[2] * e2[1], ny = e1[2] * e2[0] - e1[0] * e2[2], nz = e1[0] * e2[1] - e1[1] * e2[0];
        geo[3 * i] = make_float4(A[0], A[1], A[2], nx);
        geo[3 * i + 1] = make_float4(e1[0], e1[1], e1[2], ny);
        geo[3 * i + 2] = make_float4(e2[0], e2[1], e2[2], nz);
    }
    float4 *dgeo, *q1, *q2, *q3;
    unsigned *out;
    const int blocks = sms * 6, threads = 128;
    cudaMalloc(&dgeo, sizeof(float4) * 3 * n);
    cudaMalloc(&q1, sizeof(float4) * (path_q_f4(1, n) + 16));
    cudaMalloc(&q2, sizeof(float4) * (path_q_f4(3, n) + 16));
    cudaMalloc(&out, sizeof(unsigned) * blocks * threads);
    cudaMalloc(&q3, sizeof(float4) * (path_q_f4(5, n) + 16));
    cudaMemcpy(dgeo, geo.data(), sizeof(float4) * 3 * n, cudaMemcpyHostToDevice);
    quick_records_kernel<<<(n + 127) / 128, 128>>>(dgeo, n, 4.0f, q1, 1);
    quick_records_kernel<<<(n + 127) / 128, 128>>>(dgeo, n, 4.0f, q2, 2);
    quick_records_kernel<<<(n + 127) / 128, 128>>>(dgeo, n, 4.0f, q3, 3);
    cudaDeviceSynchronize();
    cudaEvent_t e0, e1;
    cudaEventCreate(&e0);
    cudaEventCreate(&e1);
    std::vector<unsigned> h(blocks * threads);
    unsigned long long sums[7] = {0, 0, 0, 0, 0, 0, 0};
    const char *names[7] = {"scalar FFMA, LDS.128 records", "packed FFMA2, LDS.128 pair records",
                            "packed FFMA2, __constant__ pair records (uniform operand)",
                            "scalar FFMA, __constant__ records (uniform operand)",
                            "packed FFMA2 ordered by ray component (operand reuse), LDS.128",
                            "packed FFMA2, TWO rays per lane share each LDS.128 (per ray)",
                            "lane = triangle, packed FFMA2 over two rays (scratch)"};
    for (int v = 0; v < 7; ++v) {
        const bool pk = v == 1 || v == 2 || v == 4 || v == 5;
        const size_t qf4 = path_q_f4(v == 6 ? 5 : (pk ? 3 : 1), n);
        if (v == 2 || v == 3) {
            if (qf4 > CQ_F4) {
                printf("%-62s skipped (records exceed the constant space)\n", names[v]);
                continue;
            }
            cudaMemcpyToSymbol(cq, pk ? q2 : q1, sizeof(float4) * qf4, 0, cudaMemcpyDeviceToDevice);
        }
        const size_t smem = (v < 2 || v >= 4) ? sizeof(float4) * (qf4 + (v == 6 ? PATH_SCRATCH_F4 * 4 : 0)) : 0;
        auto launch = [&]() {
            if (v == 0) probe<0><<<blocks, threads, smem>>>(q1, n, reps, out);
            if (v == 1) probe<1><<<blocks, threads, smem>>>(q2, n, reps, out);
            if (v == 2) probe<2><<<blocks, threads, smem>>>(q2, n, reps, out);
            if (v == 3) probe<3><<<blocks, threads, smem>>>(q1, n, reps, out);
            if (v == 4) probe<4><<<blocks, threads, smem>>>(q2, n, reps, out);
            if (v == 5) probe<5><<<blocks, threads, smem>>>(q2, n, reps, out);
            if (v == 6) probe<6><<<blocks, threads, smem>>>(q3, n, reps, out);
        };
        if (v == 0) cudaFuncSetAttribute(probe<0>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (v == 1) cudaFuncSetAttribute(probe<1>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (v == 4) cudaFuncSetAttribute(probe<4>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (v == 5) cudaFuncSetAttribute(probe<5>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        if (v == 6) cudaFuncSetAttribute(probe<6>, cudaFuncAttributeMaxDynamicSharedMemorySize, (int)smem);
        float best = 1e30f;
        for (int rep = 0; rep < 4; ++rep) {
            cudaEventRecord(e0);
            launch();
            cudaEventRecord(e1);
            cudaEventSynchronize(e1);
            float ms;
            cudaEventElapsedTime(&ms, e0, e1);
            if (rep && ms < best)
                best = ms;
        }
        cudaError_t err = cudaGetLastError();
        if (err != cudaSuccess) {
            printf("%-62s CUDA error %s\n", names[v], cudaGetErrorString(err));
            continue;
        }
        cudaMemcpy(h.data(), out, sizeof(unsigned) * h.size(), cudaMemcpyDeviceToHost);
        for (unsigned x : h)
            sums[v] += x;
        const int nn = v == 6 ? n : (n & ~3); // triangles the unrolled loops cover
        const double pairs = (double)blocks * threads * reps * nn * (v == 5 ? 2 : 1);
        const double clk_per_pair_warp = best * 1e-3 * khz * 1e3 * sms * 4 / (pairs / 32);
        printf("%-62s %8.3f ms  %.3e pairs/s  %5.2f SMSP-clocks per (warp, triangle)  survivors %llu\n", names[v], best,
               pairs / (best * 1e-3), clk_per_pair_warp, sums[v]);
    }
    return 0;
}
