// What-if micro-benchmark of the force inner loop alone (no TMA, no epilogue): the tile is
// loaded into shared memory once and swept `iters` times. Build variants with -DNB_UBENCH_NO_MUFU.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../raylib-nbody_b200/csrc/force.cuh"
using namespace nb;
template <int R, bool PACKED, int CH, bool FLUSH>
__global__ void __launch_bounds__(128, 4) inner(const float4* tile_g, double* out, int iters, float eps2) {
    __shared__ alignas(128) float4 tile[kTileBodies];
    for (int k = threadIdx.x; k < kTileBodies; k += 128) tile[k] = tile_g[k];
    __syncthreads();
    Target<PACKED, false> tg[R]; long long ti[R]; double S[R][4];
    for (int r = 0; r < R; ++r) { tg[r].init(0.01f * (threadIdx.x + r), 0.02f * r, 0.5f); ti[r] = -1; S[r][0] = S[r][1] = S[r][2] = S[r][3] = 0; }
    for (int it = 0; it < iters; ++it) {
#pragma unroll 1
        for (int c = 0; c < kTileBodies / CH; ++c) {
            chunk_interact<R, PACKED, false, false, CH>(tg, &tile[c * CH], eps2, 0, ti);
            if (FLUSH) for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
        }
    }
    if (!FLUSH) for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
    double a = 0; for (int r = 0; r < R; ++r) a += S[r][0] + S[r][1] + S[r][2];
    if (a == 1.2345) out[0] = a;
}
template <int R, bool PACKED, int CH, bool FLUSH> void run(const char* name, int sms, const float4* t, double* o) {
    const int iters = 400; cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    inner<R,PACKED,CH,FLUSH><<<sms * 4, 128>>>(t, o, 4, 1e-6f); cudaDeviceSynchronize();
    cudaEventRecord(e0); inner<R,PACKED,CH,FLUSH><<<sms * 4, 128>>>(t, o, iters, 1e-6f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double inter = (double)sms * 4 * 128 * R * kTileBodies * iters;
    double per_clk_sm = inter / (ms * 1e-3 * khz * 1e3) / sms;
    printf("%-34s %8.3f ms  %.3f inter/clk/SM  = %.1f%% of 20-flop peak\n", name, ms, per_clk_sm, per_clk_sm / 12.8 * 100);
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float4* h = new float4[kTileBodies];
    for (int k = 0; k < kTileBodies; ++k) h[k] = make_float4(0.001f * k, 1.f + 0.002f * k, 0.3f * (k % 7), 0.5f + 0.001f * (k % 13));
    float4* t; cudaMalloc(&t, sizeof(float4) * kTileBodies); cudaMemcpy(t, h, sizeof(float4) * kTileBodies, cudaMemcpyHostToDevice);
    double* o; cudaMalloc(&o, 8);
#ifdef NB_UBENCH_NO_MUFU
    printf("--- MUFU replaced by 2 integer ops (what-if) ---\n");
#endif
    run<4, true, 32, true>("packed R=4 chunk 32 flush", sms, t, o);
    run<4, true, 64, true>("packed R=4 chunk 64 flush", sms, t, o);
    run<4, true, 32, false>("packed R=4 chunk 32 NO flush", sms, t, o);
    run<2, true, 32, true>("packed R=2 chunk 32 flush", sms, t, o);
    run<2, true, 32, false>("packed R=2 chunk 32 NO flush", sms, t, o);
    run<4, false, 32, true>("scalar R=4 chunk 32 flush", sms, t, o);
    run<4, false, 32, false>("scalar R=4 chunk 32 NO flush", sms, t, o);
    return 0;
}
