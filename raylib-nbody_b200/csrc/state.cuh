// state.cuh — O(N) kernels around the force pass: extent/pack (FP64 state -> FP32 j-stream),
// the stand-alone integrator moves used when a call cannot ride in a force epilogue, and
// zero_net_momentum. All reductions are two-level with a fixed-order final pass.
#pragma once
#include "epilogue.cuh"

namespace nb {

constexpr int kStateThreads = 256;

struct StateView {
    double* x[3];
    double* v[3];
    double* a[3];
    double* a0[3];
    const float* mass;
    const uint8_t* pinned;
    uint8_t* live;
    long long n;
    long long lo, hi;  // rows this rank owns
};

// max |coordinate| over bodies passing the gather filter -> extent slots `slot_a`, `slot_b`.
__global__ void __launch_bounds__(kStateThreads) extent_kernel(StateView s, unsigned int* extent, int slot_a,
                                                               int slot_b) {
    __shared__ float red[kStateThreads / 32];
    float e = 0.f;
    for (long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < s.hi;
         i += (long long)gridDim.x * blockDim.x) {
        const double x[3] = {s.x[0][i], s.x[1][i], s.x[2][i]};
        const double v[3] = {s.v[0][i], s.v[1][i], s.v[2][i]};
        e = fmaxf(e, extent_of(x, body_live(x, v, s.mass[i])));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e = fmaxf(e, __shfl_xor_sync(0xffffffffu, e, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = e;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kStateThreads / 32; ++w) e = fmaxf(e, red[w]);
        atomicMax(&extent[slot_a], __float_as_uint(e));
        atomicMax(&extent[slot_b], __float_as_uint(e));
    }
}

// FP64 state -> pair-SoA FP32 j-stream of generation `gen` (+ live flags), rows [lo,hi).
__global__ void __launch_bounds__(kStateThreads) pack_kernel(StateView s, float* pk, const unsigned int* extent,
                                                             int gen, int mass_exp, double g, float softening) {
    const Scale sc = scale_for(extent[(gen + kExtentSlots - 1) % kExtentSlots], mass_exp, g, softening);
    const double inv_mass = pow2d(-mass_exp);
    const long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= s.hi) return;
    const double x[3] = {s.x[0][i], s.x[1][i], s.x[2][i]};
    const double v[3] = {s.v[0][i], s.v[1][i], s.v[2][i]};
    const float m = s.mass[i];
    const bool live = body_live(x, v, m);
    s.live[i] = live ? 1 : 0;
    pack_body(pk, i, x, m, live, sc.inv_len, inv_mass);
}

// Velocity Verlet drift of the first substep of a call, Physics.hpp:297-304.
__global__ void __launch_bounds__(kStateThreads) verlet_drift_kernel(StateView s, double dt, double half_dt2) {
    const long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= s.hi || s.pinned[i]) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double a = s.a[k][i];
        s.x[k][i] = __dadd_rn(s.x[k][i], __dadd_rn(__dmul_rn(s.v[k][i], dt), __dmul_rn(a, half_dt2)));
        s.a0[k][i] = a;
    }
}

// Semi-implicit Euler kick+drift with the stored acceleration, Physics.hpp:276-290.
__global__ void __launch_bounds__(kStateThreads) euler_move_kernel(StateView s, double dt, float max_speed) {
    const long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= s.hi || s.pinned[i]) return;
    double v[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) v[k] = __dadd_rn(s.v[k][i], __dmul_rn(s.a[k][i], dt));
    cap_speed(v, max_speed);
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        s.v[k][i] = v[k];
        s.x[k][i] = __dadd_rn(s.x[k][i], __dmul_rn(v[k], dt));
    }
}

// ---- zero_net_momentum, Physics.hpp:66-85 ------------------------------------------------
// pass 1: pinned -> v = 0; per-CTA sums of m*v and m over non-pinned rows [lo,hi)
__global__ void __launch_bounds__(kStateThreads) momentum_partial_kernel(StateView s, double* part /*[grid][4]*/) {
    __shared__ double red[kStateThreads / 32][4];
    double acc[4] = {0.0, 0.0, 0.0, 0.0};
    for (long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x; i < s.hi;
         i += (long long)gridDim.x * blockDim.x) {
        if (s.pinned[i]) {
            s.v[0][i] = 0.0; s.v[1][i] = 0.0; s.v[2][i] = 0.0;
            continue;
        }
        const double m = (double)s.mass[i];
#pragma unroll
        for (int k = 0; k < 3; ++k) acc[k] += __dmul_rn(m, s.v[k][i]);
        acc[3] += m;
    }
#pragma unroll
    for (int k = 0; k < 4; ++k) {
        double w = acc[k];
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) w += __shfl_xor_sync(0xffffffffu, w, o);
        if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5][k] = w;
    }
    __syncthreads();
    if (threadIdx.x < 4) {
        double w = 0.0;
        for (int q = 0; q < kStateThreads / 32; ++q) w += red[q][threadIdx.x];
        part[blockIdx.x * 4 + threadIdx.x] = w;
    }
}

// fixed-order final pass: out[k] = sum_b part[b][k], one thread per value
__global__ void sum_rows_kernel(const double* part, int rows, int width, int stride, double* out) {
    const int k = threadIdx.x;
    if (k >= width) return;
    double acc = 0.0;
    for (int b = 0; b < rows; ++b) acc += part[(long long)b * stride + k];
    out[k] = acc;
}

// pass 2: v -= P/M on non-pinned rows; totals[0..2] = P, totals[3] = M (global sums)
__global__ void __launch_bounds__(kStateThreads) momentum_apply_kernel(StateView s, const double* totals) {
    const double mt = totals[3];
    if (!(mt > 0.0)) return;  // Physics.hpp:77
    const long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= s.hi || s.pinned[i]) return;
#pragma unroll
    for (int k = 0; k < 3; ++k) s.v[k][i] = __dadd_rn(s.v[k][i], -__ddiv_rn(totals[k], mt));
}

}  // namespace nb
