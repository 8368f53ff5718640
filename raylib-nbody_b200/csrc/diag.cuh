// diag.cuh — stand-alone Physics::compute_diagnostics (src/systems/Physics.hpp:117-177):
// KE, PE, E, P, COM, M over ALL bodies, unfiltered (unlike the gravity gather filter).
// PE = -G sum_{i<j} m_i m_j / sqrt(r^2+eps^2) is evaluated as -G/2 sum_i m_i phi_i with
// phi_i = sum_{j != i} m_j / r_ij by the potential-only instantiation of the force kernel
// (force_kernel<..., PONLY>: TMA ring, pair records, R targets per thread, 7 FMA-pipe instructions
// per source pair) over a stream packed with this function's own rules (pack_kernel<true>: every body
// with a finite position, raw masses); every other sum is FP64 in that kernel's epilogue
// (diag_accumulate). Warp-shuffle + block reduction per CTA, then a fixed-order final pass
// (sum_rows_kernel), so results are bit-reproducible. This file keeps the extent pass and the FP64
// kernel of the exact path.
#pragma once
#include "epilogue.cuh"
#include "ptx.cuh"
#include "state.cuh"

namespace nb {

constexpr int kDiagThreads = 256;
constexpr int kDiagExtentSlot = kExtentSlots;  // extra slot after the rotating ones

__global__ void __launch_bounds__(kDiagThreads) diag_extent_kernel(const double* x0, const double* x1,
                                                                   const double* x2, long long lo, long long hi,
                                                                   PeerSet out) {
    __shared__ float red[kDiagThreads / 32];
    float e = 0.f;
    for (long long i = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < hi;
         i += (long long)gridDim.x * blockDim.x) {
        const double x[3] = {x0[i], x1[i], x2[i]};
        e = fmaxf(e, extent_of(x, finite_d(x[0]) && finite_d(x[1]) && finite_d(x[2])));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e = fmaxf(e, __shfl_xor_sync(0xffffffffu, e, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = e;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kDiagThreads / 32; ++w) e = fmaxf(e, red[w]);
        for (int q = 0; q < out.n; ++q) atomicMax(&out.extent[q][kDiagExtentSlot], __float_as_uint(e));
    }
}

// FP64 variant for the exact path: the reference's pair terms -G*m_i*m_j/sqrt(r2) over i < j
// (Physics.hpp:148-154) in FP64, one target per thread; only the summation order differs.
__global__ void __launch_bounds__(kDiagThreads) diag_exact_kernel(StateView s, double g, float softening,
                                                                  double* part) {
    __shared__ double sx[kDiagThreads], sy[kDiagThreads], sz[kDiagThreads], sm[kDiagThreads];
    __shared__ double red[kDiagThreads / 32][kDiagValues];
    const double eps2 = (double)softening * (double)softening;
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const bool valid = i < s.n;
    double xi[3] = {0, 0, 0}, vi[3] = {0, 0, 0};
    float mi = 0.f;
    if (valid) {
        for (int k = 0; k < 3; ++k) { xi[k] = s.x[k][i]; vi[k] = s.v[k][i]; }
        mi = s.mass[i];
    }
    double pe = 0.0;
    const long long j_start = (blockIdx.x * (long long)blockDim.x / kDiagThreads) * kDiagThreads;
    for (long long j0 = j_start; j0 < s.n; j0 += kDiagThreads) {
        const long long j = j0 + threadIdx.x;
        if (j < s.n) {
            sx[threadIdx.x] = s.x[0][j]; sy[threadIdx.x] = s.x[1][j]; sz[threadIdx.x] = s.x[2][j];
            sm[threadIdx.x] = (double)s.mass[j];
        }
        __syncthreads();
        if (valid) {
            const int cnt = (int)min((long long)kDiagThreads, s.n - j0);
            for (int u = 0; u < cnt; ++u) {
                if (j0 + u <= i) continue;
                const double dx = sx[u] - xi[0], dy = sy[u] - xi[1], dz = sz[u] - xi[2];
                const double r = __dsqrt_rn(__dadd_rn(__dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), __dmul_rn(dz, dz)), eps2));
                pe += __ddiv_rn(__dmul_rn(__dmul_rn(-g, (double)mi), sm[u]), r);
            }
        }
        __syncthreads();
    }
    DiagTerms dg;
#pragma unroll
    for (int k = 0; k < kDiagValues; ++k) dg.v[k] = 0.0;
    if (valid) {
        diag_accumulate(dg, xi, vi, mi, 0.0, g, false);
        dg.v[1] = pe;
    }
#pragma unroll
    for (int k = 0; k < 10; ++k) {
        double w = dg.v[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = w;
    }
    __syncthreads();
    if (threadIdx.x < 10) {
        double w = 0.0;
        for (int q = 0; q < kDiagThreads / 32; ++q) w += red[q][threadIdx.x];
        part[(long long)blockIdx.x * kDiagValues + threadIdx.x] = w;
    }
}

}  // namespace nb
