// Micro-benchmark: do the FP64 pipe and the FP32 (FFMA2) pipe of a B200 SM run concurrently?
// A CTA has NW32 warps sweeping a shared-memory tile with the packed FP32 interaction and NW64
// warps sweeping it with an FP64 interaction (MUFU.RSQ64H seed, no Newton step). Times:
// FP32 warps alone, FP64 warps alone, both. Also prints the accuracy of rsqrt.approx.f64.
#include <cstdio>
#include <cmath>
#include <cuda_runtime.h>
#include "../../raylib-nbody_b200/csrc/force.cuh"
using namespace nb;

__device__ __forceinline__ double rsqrt64h(double x) {
    double y;
    asm("rsqrt.approx.ftz.f64 %0, %1;" : "=d"(y) : "d"(x));
    return y;
}

constexpr int kT = 512;

template <int R>
__device__ __forceinline__ void sweep64(const double* xs, const double* ys, const double* zs, const double* ms,
                                        int j0, int j1, const double (&xi)[R], const double (&yi)[R],
                                        const double (&zi)[R], double (&ax)[R], double (&ay)[R], double (&az)[R],
                                        double eps2) {
#pragma unroll 2
    for (int j = j0; j < j1; ++j) {
        const double xj = xs[j], yj = ys[j], zj = zs[j], mj = ms[j];
#pragma unroll
        for (int r = 0; r < R; ++r) {
            const double dx = xj - xi[r], dy = yj - yi[r], dz = zj - zi[r];
            double r2 = fma(dx, dx, eps2);
            r2 = fma(dy, dy, r2);
            r2 = fma(dz, dz, r2);
            const double y = rsqrt64h(r2);
            const double s = (y * y) * (y * mj);
            ax[r] = fma(s, dx, ax[r]);
            ay[r] = fma(s, dy, ay[r]);
            az[r] = fma(s, dz, az[r]);
        }
    }
}

// warps [0, NW32): FP32 packed sweep of sources [0, n32); warps [NW32, NW32+NW64): FP64 sweep of [512-n64, 512)
template <int NW32, int NW64, int R, int MINB>
__global__ void __launch_bounds__((NW32 + NW64) * 32, MINB)
mix(const float4* tile_g, double* out, int iters, int n32, int n64, float eps2) {
    __shared__ alignas(128) float4 tile[kT];
    __shared__ alignas(16) double xs[kT], ys[kT], zs[kT], ms[kT];
    for (int k = threadIdx.x; k < kT; k += blockDim.x) {
        tile[k] = tile_g[k];
        const float* f = reinterpret_cast<const float*>(tile_g) + (k >> 1) * 8 + (k & 1);
        xs[k] = f[0]; ys[k] = f[2]; zs[k] = f[4]; ms[k] = f[6];
    }
    __syncthreads();
    const int warp = threadIdx.x >> 5;
    double acc = 0;
    if (warp < NW32) {
        Target<true, false> tg[R]; long long ti[R]; double S[R][4];
        for (int r = 0; r < R; ++r) { tg[r].init(0.01f * (threadIdx.x + r), 0.02f * r, 0.5f); ti[r] = -1; S[r][0] = S[r][1] = S[r][2] = S[r][3] = 0; }
        for (int it = 0; it < iters; ++it) {
#pragma unroll 1
            for (int c = 0; c < n32 / 64; ++c) {
                chunk_interact<R, true, false, false, 64>(tg, &tile[c * 64], eps2, 0, ti);
                for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
            }
        }
        for (int r = 0; r < R; ++r) acc += S[r][0] + S[r][1] + S[r][2];
    } else {
        double xi[R], yi[R], zi[R], ax[R], ay[R], az[R];
        for (int r = 0; r < R; ++r) { xi[r] = 0.01 * (threadIdx.x + r); yi[r] = 0.02 * r; zi[r] = 0.5; ax[r] = ay[r] = az[r] = 0; }
        for (int it = 0; it < iters; ++it) sweep64<R>(xs, ys, zs, ms, kT - n64, kT, xi, yi, zi, ax, ay, az, (double)eps2);
        for (int r = 0; r < R; ++r) acc += ax[r] + ay[r] + az[r];
    }
    if (acc == 1.2345) out[0] = acc;
}

template <int NW32, int NW64, int R, int MINB>
void run(const char* name, int sms, const float4* t, double* o, int n32, int n64) {
    const int iters = 300; cudaEvent_t e0, e1; cudaEventCreate(&e0); cudaEventCreate(&e1);
    auto fn = mix<NW32, NW64, R, MINB>;
    int per_sm = 0; cudaOccupancyMaxActiveBlocksPerMultiprocessor(&per_sm, fn, (NW32 + NW64) * 32, 0);
    fn<<<sms * per_sm, (NW32 + NW64) * 32>>>(t, o, 4, n32, n64, 1e-6f); cudaDeviceSynchronize();
    cudaEventRecord(e0); fn<<<sms * per_sm, (NW32 + NW64) * 32>>>(t, o, iters, n32, n64, 1e-6f); cudaEventRecord(e1); cudaEventSynchronize(e1);
    float ms; cudaEventElapsedTime(&ms, e0, e1);
    cudaError_t err = cudaGetLastError();
    int khz; cudaDeviceGetAttribute(&khz, cudaDevAttrClockRate, 0);
    double inter = (double)sms * per_sm * 32 * R * iters * ((double)NW32 * n32 + (double)NW64 * n64);
    double per_clk_sm = inter / (ms * 1e-3 * khz * 1e3) / sms;
    printf("%-30s ctas/sm %d n32 %3d n64 %3d %8.3f ms  %.3f inter/clk/SM = %.1f%% of 20-flop peak %s\n", name, per_sm, n32, n64, ms,
           per_clk_sm, per_clk_sm / 12.8 * 100, err == cudaSuccess ? "" : cudaGetErrorString(err));
}

__global__ void rsq_err(double* out) {
    double worst = 0;
    for (int k = threadIdx.x; k < (1 << 20); k += blockDim.x) {
        double x = 1.0 + 3.0 * (k + 0.37) / (1 << 20);  // [1,4): one full period of rsqrt
        double y = rsqrt64h(x), ref = 1.0 / sqrt(x);
        worst = fmax(worst, fabs(y - ref) / ref);
    }
    for (int o = 16; o > 0; o >>= 1) worst = fmax(worst, __shfl_xor_sync(0xffffffffu, worst, o));
    if (threadIdx.x == 0) out[0] = worst;
}

int main() {
    int sms; cudaDeviceGetAttribute(&sms, cudaDevAttrMultiProcessorCount, 0);
    float4* h = new float4[kT];
    for (int k = 0; k < kT; ++k) h[k] = make_float4(0.001f * k, 1.f + 0.002f * k, 0.3f * (k % 7), 0.5f + 0.001f * (k % 13));
    float4* t; cudaMalloc(&t, sizeof(float4) * kT); cudaMemcpy(t, h, sizeof(float4) * kT, cudaMemcpyHostToDevice);
    double* o; cudaMalloc(&o, 8);
    rsq_err<<<1, 32>>>(o); double e; cudaMemcpy(&e, o, 8, cudaMemcpyDeviceToHost);
    printf("rsqrt.approx.ftz.f64 max rel err on [1,4): %.3e (2^%.1f)\n", e, log2(e));
    run<4, 0, 4, 4>("fp32 only 4w x4", sms, t, o, 512, 0);
    run<4, 4, 4, 2>("fp32 only (4+4 idle) x2", sms, t, o, 512, 0);
    run<4, 4, 4, 2>("fp64 only (4 idle+4) x2", sms, t, o, 0, 512);
    run<0, 4, 4, 4>("fp64 only 4w x4", sms, t, o, 0, 512);
    run<4, 4, 4, 2>("mix 4+4 x2  384/128", sms, t, o, 384, 128);
    run<4, 4, 4, 2>("mix 4+4 x2  320/192", sms, t, o, 320, 192);
    run<4, 4, 4, 2>("mix 4+4 x2  384/192", sms, t, o, 384, 192);
    run<4, 4, 4, 2>("mix 4+4 x2  448/ 64", sms, t, o, 448, 64);
    run<8, 4, 4, 1>("mix 8+4 x1  448/128", sms, t, o, 448, 128);
    run<8, 4, 4, 1>("mix 8+4 x1  384/256", sms, t, o, 384, 256);
    run<8, 4, 4, 1>("mix 8+4 x1  512/  0", sms, t, o, 512, 0);
    run<8, 8, 4, 1>("mix 8+8 x1  384/128", sms, t, o, 384, 128);
    run<8, 8, 4, 1>("mix 8+8 x1  384/192", sms, t, o, 384, 192);
    run<8, 8, 2, 1>("mix 8+8 x1 R2 384/128", sms, t, o, 384, 128);
    run<12, 4, 4, 1>("mix 12+4 x1 (85 regs) 448/192", sms, t, o, 448, 192);
    run<12, 4, 4, 1>("mix 12+4 x1 (85 regs) 512/  0", sms, t, o, 512, 0);
    return 0;
}
