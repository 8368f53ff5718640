// What-if micro-benchmark of the packed inner loop: which ingredient keeps the FMA pipe at ~88 %
// instead of ~97 %? Flags: MUFU (real rsqrt vs a free register copy), LDS (sources from shared
// memory vs registers), R (targets per thread), and the CTA count per SM.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../raylib-nbody_b200/csrc/ptx.cuh"
using namespace nb;
constexpr int kT = 512;

template <int R, int MUFU, bool LDS, int MINB, bool EQM = false>
__global__ void __launch_bounds__(128, MINB) inner(const float4* tile_g, float* out, int iters, float eps2f) {
    __shared__ alignas(128) float4 tile[kT];
    for (int k = threadIdx.x; k < kT; k += 128) tile[k] = tile_g[k];
    __syncthreads();
    f2 nx[R], ny[R], nz[R], ax[R], ay[R], az[R];
    for (int r = 0; r < R; ++r) {
        const float x = 0.01f * (threadIdx.x + r);
        nx[r] = f2_make(-x, -x); ny[r] = f2_make(-0.02f * r, -0.02f * r); nz[r] = f2_make(-0.5f, -0.5f);
        ax[r] = ay[r] = az[r] = f2_make(0.f, 0.f);
    }
    const f2 e2 = f2_make(eps2f, eps2f);
    ulonglong2 A0 = *reinterpret_cast<const ulonglong2*>(&tile[threadIdx.x & 3]);
    ulonglong2 B0 = *reinterpret_cast<const ulonglong2*>(&tile[4 + (threadIdx.x & 3)]);
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int c = 0; c < kT / 64; ++c) {
#pragma unroll
            for (int u = 0; u < 32; ++u) {
                ulonglong2 A, B;
                if (LDS) {
                    A = *reinterpret_cast<const ulonglong2*>(&tile[c * 64 + 2 * u]);
                    B = *reinterpret_cast<const ulonglong2*>(&tile[c * 64 + 2 * u + 1]);
                } else {
                    asm volatile("" : "+l"(A0.x), "+l"(A0.y), "+l"(B0.x), "+l"(B0.y));
                    A = A0; B = B0;
                }
                const f2 xj{A.x}, yj{A.y}, zj{B.x}, mj{B.y};
#pragma unroll
                for (int r = 0; r < R; ++r) {
                    const f2 dx = f2_add(xj, nx[r]), dy = f2_add(yj, ny[r]), dz = f2_add(zj, nz[r]);
                    f2 r2 = f2_fma(dx, dx, e2);
                    r2 = f2_fma(dy, dy, r2);
                    r2 = f2_fma(dz, dz, r2);
                    const f2 r6 = f2_mul(f2_mul(r2, r2), r2);
                    f2 w = r6;
                    if (MUFU == 1) { float lo, hi; f2_split(r6, lo, hi); w = f2_make(rsqrt_fast(lo), rsqrt_fast(hi)); }
                    if (MUFU == 2) { float lo, hi; f2_split(r6, lo, hi); w = f2_make(rsqrt_fast(lo), hi); }
                    const f2 s = EQM ? w : f2_mul(mj, w);
                    ax[r] = f2_fma(s, dx, ax[r]);
                    ay[r] = f2_fma(s, dy, ay[r]);
                    az[r] = f2_fma(s, dz, az[r]);
                }
            }
        }
    }
    float a = 0;
    for (int r = 0; r < R; ++r) { float lo, hi; f2_split(ax[r], lo, hi); a += lo + hi; f2_split(ay[r], lo, hi); a += lo + hi; f2_split(az[r], lo, hi); a += lo + hi; }
    if (a == 1.2345f) out[0] = a;
}
template <int R, int MUFU, bool LDS, int MINB, bool EQM = false> void run(const char* name, int sms, const float4* t, float* o) {
    const int iters = 300; cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto fn = inner<R, MUFU, LDS, MINB, EQM>;
    int per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, 128, 0);
    fn<<<sms * per_sm, 128>>>(t, o, 4, 1e-6f); cudaDeviceSynchronize();
    cudaEventRecord(e0); fn<<<sms * per_sm, 128>>>(t, o, iters, 1e-6f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double inter = (double)sms * per_sm * 128 * R * kT * iters;
    double per_clk_sm = inter / (ms * 1e-3 * khz * 1e3) / sms;
    printf("%-34s ctas/sm %d %8.3f ms  %.3f inter/clk/SM  = %.1f%% of 20-flop peak, FMA pipe %.1f%%\n", name, per_sm, ms, per_clk_sm,
           per_clk_sm / 12.8 * 100, per_clk_sm / (EQM ? 11.6364 : 10.6667) * 100);
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float4* h = new float4[kT];
    for (int k = 0; k < kT; ++k) h[k] = make_float4(0.001f * k, 1.f + 0.002f * k, 0.3f * (k % 7), 0.5f + 0.001f * (k % 13));
    float4* t; cudaMalloc(&t, sizeof(float4) * kT); cudaMemcpy(t, h, sizeof(float4) * kT, cudaMemcpyHostToDevice);
    float* o; cudaMalloc(&o, 8);
    run<4, 1, true, 4>("R4 mufu lds x4", sms, t, o);
    run<4, 0, true, 4>("R4 NOmufu lds x4", sms, t, o);
    run<4, 2, true, 4>("R4 HALFmufu lds x4", sms, t, o);
    run<4, 1, true, 4, true>("R4 mufu lds x4 EQUAL-MASS(11 ops)", sms, t, o);
    run<4, 0, true, 4, true>("R4 NOmufu lds x4 EQUAL-MASS", sms, t, o);
    run<4, 1, true, 5>("R4 mufu lds x5", sms, t, o);
    run<4, 1, true, 5, true>("R4 mufu lds x5 EQUAL-MASS", sms, t, o);
    run<2, 1, true, 8>("R2 mufu lds x8", sms, t, o);
    run<2, 0, true, 8>("R2 NOmufu lds x8", sms, t, o);
    run<3, 1, true, 6>("R3 mufu lds x6", sms, t, o);
    run<3, 1, true, 5>("R3 mufu lds x5", sms, t, o);
    return 0;
}
