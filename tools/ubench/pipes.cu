// Micro-benchmark: throughput of FFMA2 vs FFMA and whether they overlap (fmaheavy / fmalite).
#include <cstdio>
#include <cuda_runtime.h>
struct f2 { unsigned long long r; };
__device__ __forceinline__ f2 mk(float a, float b){ f2 o; asm("mov.b64 %0, {%1,%2};":"=l"(o.r):"f"(a),"f"(b)); return o; }
__device__ __forceinline__ f2 fma2(f2 a, f2 b, f2 c){ f2 o; asm volatile("fma.rn.f32x2 %0,%1,%2,%3;":"=l"(o.r):"l"(a.r),"l"(b.r),"l"(c.r)); return o; }
__device__ __forceinline__ float fma1(float a, float b, float c){ float o; asm volatile("fma.rn.f32 %0,%1,%2,%3;":"=f"(o):"f"(a),"f"(b),"f"(c)); return o; }
__device__ __forceinline__ float rsq(float a){ float o; asm volatile("rsqrt.approx.ftz.f32 %0,%1;":"=f"(o):"f"(a)); return o; }

template<int NP, int NS, int NM>
__global__ void __launch_bounds__(256) k(float* out, int iters, float seed) {
    f2 p[NP > 0 ? NP : 1]; float s[NS > 0 ? NS : 1]; float m[NM > 0 ? NM : 1];
    for (int i = 0; i < NP; ++i) p[i] = mk(seed + i, seed - i);
    for (int i = 0; i < NS; ++i) s[i] = seed + 0.5f * i;
    for (int i = 0; i < NM; ++i) m[i] = seed + 2.f + i;
    const f2 a2 = mk(1.0001f, 0.9999f), b2 = mk(1e-7f, -1e-7f);
    for (int it = 0; it < iters; ++it) {
#pragma unroll
        for (int u = 0; u < 8; ++u) {
#pragma unroll
            for (int i = 0; i < NP; ++i) p[i] = fma2(p[i], a2, b2);
#pragma unroll
            for (int i = 0; i < NS; ++i) s[i] = fma1(s[i], 1.0001f, 1e-7f);
#pragma unroll
            for (int i = 0; i < NM; ++i) m[i] = rsq(m[i]);
        }
    }
    float acc = 0;
    for (int i = 0; i < NP; ++i) { float lo, hi; asm("mov.b64 {%0,%1}, %2;":"=f"(lo),"=f"(hi):"l"(p[i].r)); acc += lo + hi; }
    for (int i = 0; i < NS; ++i) acc += s[i];
    for (int i = 0; i < NM; ++i) acc += m[i];
    if (acc == 12345.678f) out[0] = acc;
}
template<int NP, int NS, int NM> void run(const char* name, int sms, float* d) {
    const int iters = 20000; cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<NP,NS,NM><<<sms*4, 256>>>(d, 100, 1.f); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<NP,NS,NM><<<sms*4, 256>>>(d, iters, 1.f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double cyc = ms * 1e-3 * khz * 1e3;               // at max clock
    double warps_per_smsp = 4.0 * 8 / 4;                // 4 blocks x 8 warps over 4 SMSPs
    double inst = (double)iters * 8 * (NP + NS + NM) * warps_per_smsp;
    printf("%-28s %8.3f ms  issue/clk/SMSP %.3f  FMA lanes/clk/SM %.1f  MUFU lanes/clk/SM %.1f\n", name, ms, inst / cyc,
           (double)iters * 8 * (2.0 * NP + NS) * warps_per_smsp * 32 * 4 / cyc, (double)iters * 8 * NM * warps_per_smsp * 32 * 4 / cyc);
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float* d; cudaMalloc(&d, 4);
    run<8,0,0>("FFMA2 x8", sms, d);
    run<0,8,0>("FFMA x8", sms, d);
    run<0,16,0>("FFMA x16", sms, d);
    run<8,8,0>("FFMA2 x8 + FFMA x8", sms, d);
    run<8,4,0>("FFMA2 x8 + FFMA x4", sms, d);
    run<8,2,0>("FFMA2 x8 + FFMA x2", sms, d);
    run<0,0,8>("MUFU x8", sms, d);
    run<8,0,2>("FFMA2 x8 + MUFU x2", sms, d);
    run<12,0,2>("FFMA2 x12 + MUFU x2", sms, d);
    run<12,0,4>("FFMA2 x12 + MUFU x4", sms, d);
    run<0,12,1>("FFMA x12 + MUFU x1", sms, d);
    return 0;
}
