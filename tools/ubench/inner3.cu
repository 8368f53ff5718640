// Where do the 7 % between the bare inner loop (inner.cu) and force_kernel go? Same chunk code,
// adding one ingredient at a time: MODE 0 tile resident in shared memory; MODE 1 tiles streamed
// from global memory through the 2-stage TMA ring with a CTA barrier per tile; GRID 0 one wave of
// persistent CTAs, GRID 1 many short-lived CTAs (7 waves) like the real launch.
#include <cstdio>
#include <cuda_runtime.h>
#include "../../raylib-nbody_b200/csrc/force.cuh"
using namespace nb;
constexpr int kT = 512;

template <int R, int MODE, bool EQM>
__global__ void __launch_bounds__(128, 4) k(const float* pk, double* out, int n_tiles, float eps2) {
    __shared__ alignas(128) float4 ring[2][kT];
    __shared__ alignas(8) uint64_t full_bar[2];
    const int tid = threadIdx.x;
    Target<true, false, EQM, false> tg[R]; long long ti[R]; double S[R][4];
    for (int r = 0; r < R; ++r) { tg[r].init(0.01f * (tid + r) + blockIdx.x, 0.02f * r, 0.5f); ti[r] = -1; S[r][0] = S[r][1] = S[r][2] = S[r][3] = 0; }
    if (MODE == 0) {
        for (int q = tid; q < kT; q += 128) ring[0][q] = reinterpret_cast<const float4*>(pk)[q];
        __syncthreads();
        for (int t = 0; t < n_tiles; ++t) {
#pragma unroll 1
            for (int c = 0; c < kT / 64; ++c) {
                chunk_interact<R, true, false, false, 64, EQM, false>(tg, &ring[0][c * 64], eps2, 0, ti);
                for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
            }
        }
    } else {
        if (tid == 0) { mbar_init(&full_bar[0], 1); mbar_init(&full_bar[1], 1); fence_mbar_init(); }
        __syncthreads();
        if (tid == 0)
            for (int s = 0; s < 2 && s < n_tiles; ++s) {
                mbar_arrive_expect_tx(&full_bar[s], kT * 16);
                tma_load_1d(&ring[s][0], pk + (size_t)s * kT * 4, kT * 16, &full_bar[s]);
            }
        for (int t = 0; t < n_tiles; ++t) {
            const int s = t & 1;
            mbar_wait(&full_bar[s], (uint32_t)((t >> 1) & 1));
#pragma unroll 1
            for (int c = 0; c < kT / 64; ++c) {
                chunk_interact<R, true, false, false, 64, EQM, false>(tg, &ring[s][c * 64], eps2, 0, ti);
                for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
            }
            __syncthreads();
            if (tid == 0 && t + 2 < n_tiles) {
                mbar_arrive_expect_tx(&full_bar[s], kT * 16);
                tma_load_1d(&ring[s][0], pk + (size_t)(t + 2) * kT * 4, kT * 16, &full_bar[s]);
            }
        }
    }
    double a = 0; for (int r = 0; r < R; ++r) a += S[r][0] + S[r][1] + S[r][2];
    if (a == 1.2345) out[0] = a;
}

template <int R, int MODE, bool EQM> void run(const char* name, int sms, const float* pk, double* o, int grid, int n_tiles) {
    cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    k<R, MODE, EQM><<<grid, 128>>>(pk, o, 8, 1e-6f); cudaDeviceSynchronize();
    cudaEventRecord(e0); k<R, MODE, EQM><<<grid, 128>>>(pk, o, n_tiles, 1e-6f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    double inter = (double)grid * 128 * R * kT * n_tiles;
    double per_clk_sm = inter / (ms * 1e-3 * 1965e6) / sms;
    printf("%-44s grid %5d tiles %5d %8.3f ms  %.3f inter/clk/SM = %.1f%% of 20-flop peak %s\n", name, grid, n_tiles, ms, per_clk_sm,
           per_clk_sm / 12.8 * 100, cudaGetErrorString(cudaGetLastError()));
}
int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    const int tiles = 2048;  // 2^20 sources
    size_t nf = (size_t)tiles * kT * 4;
    float* h = new float[nf];
    for (size_t q = 0; q < nf; ++q) h[q] = 0.001f * (float)(q % 9973) + 0.5f;
    float* pk; cudaMalloc(&pk, nf * 4); cudaMemcpy(pk, h, nf * 4, cudaMemcpyHostToDevice);
    double* o; cudaMalloc(&o, 8);
    run<4, 0, false>("resident tile, one wave", sms, pk, o, sms * 4, tiles / 4);
    run<4, 1, false>("TMA ring + barrier, one wave", sms, pk, o, sms * 4, tiles / 4);
    run<4, 1, false>("TMA ring + barrier, 2048 CTAs (3.46 waves)", sms, pk, o, 2048, tiles / 4);
    run<4, 1, false>("TMA ring + barrier, 4096 CTAs x 2^19 src", sms, pk, o, 4096, tiles / 2);
    run<4, 1, false>("TMA ring + barrier, 592*7 CTAs (7 full waves)", sms, pk, o, sms * 28, tiles / 4);
    run<4, 0, true>("EQM resident tile, one wave", sms, pk, o, sms * 4, tiles / 4);
    run<4, 1, true>("EQM TMA ring + barrier, one wave", sms, pk, o, sms * 4, tiles / 4);
    run<4, 1, true>("EQM TMA ring + barrier, 592*7 CTAs", sms, pk, o, sms * 28, tiles / 4);
    return 0;
}
