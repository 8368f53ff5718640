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
    uint8_t* live;     // gather-filter flags of the current generation
    long long n;
    long long lo, hi;  // rows this rank owns
};

// max |coordinate| over bodies passing the gather filter -> extent slots `slot_a`, `slot_b`.
__global__ void __launch_bounds__(kStateThreads) extent_kernel(StateView s, PeerSet out, int slot_a, int slot_b) {
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
        for (int q = 0; q < out.n; ++q) {  // every device's copy (peer atomics over NVLink)
            atomicMax(&out.extent[q][slot_a], __float_as_uint(e));
            atomicMax(&out.extent[q][slot_b], __float_as_uint(e));
        }
    }
}

// ---- Hilbert keys: the stream is ordered along a Hilbert curve so every group of kGroupBodies rows is a
// compact clump (a Z curve's segments straddle its big jumps: measured 3-4x more "near" group visits) -------
__device__ __forceinline__ unsigned long long spread3(unsigned int v) {  // 21 bits -> every third bit
    unsigned long long x = v & 0x1fffffu;
    x = (x | x << 32) & 0x1f00000000ffffull;
    x = (x | x << 16) & 0x1f0000ff0000ffull;
    x = (x | x << 8) & 0x100f00f00f00f00full;
    x = (x | x << 4) & 0x10c30c30c30c30c3ull;
    x = (x | x << 2) & 0x1249249249249249ull;
    return x;
}
__device__ __forceinline__ unsigned long long spread2(unsigned int v) {  // 31 bits -> every second bit
    unsigned long long x = v & 0x7fffffffu;
    x = (x | x << 16) & 0x0000ffff0000ffffull;
    x = (x | x << 8) & 0x00ff00ff00ff00ffull;
    x = (x | x << 4) & 0x0f0f0f0f0f0f0f0full;
    x = (x | x << 2) & 0x3333333333333333ull;
    x = (x | x << 1) & 0x5555555555555555ull;
    return x;
}

// Skilling's axes-to-transpose (Programming the Hilbert curve, 2004): cell coordinates of `bits` bits per
// axis -> the Hilbert index, bit-interleaved with axis 0 most significant.
template <int DIMS>
__device__ __forceinline__ unsigned long long hilbert_key(unsigned int (&X)[3], int bits) {
    const unsigned int M = 1u << (bits - 1);
    for (unsigned int Q = M; Q > 1; Q >>= 1) {
        const unsigned int P = Q - 1;
#pragma unroll
        for (int i = 0; i < DIMS; ++i) {
            if (X[i] & Q) {
                X[0] ^= P;
            } else {
                const unsigned int t = (X[0] ^ X[i]) & P;
                X[0] ^= t;
                X[i] ^= t;
            }
        }
    }
#pragma unroll
    for (int i = 1; i < DIMS; ++i) X[i] ^= X[i - 1];
    unsigned int t = 0;
    for (unsigned int Q = M; Q > 1; Q >>= 1)
        if (X[DIMS - 1] & Q) t ^= Q - 1;
#pragma unroll
    for (int i = 0; i < DIMS; ++i) X[i] ^= t;
    if (DIMS == 3) return (spread3(X[0]) << 2) | (spread3(X[1]) << 1) | spread3(X[2]);
    return (spread2(X[0]) << 1) | spread2(X[1]);
}

// keys[q] / vals[q] for body row lo + q: Hilbert key (3 x 21 bits, or 2 x 31 bits for planar worlds) of the
// position inside the cube [-2^ex, 2^ex)^dims that holds every attracting body (extent slot), ~0 for rows the
// gather filter drops.
template <int DIMS>
__global__ void __launch_bounds__(kStateThreads) curve_key_kernel(StateView s, const unsigned int* extent, int slot,
                                                                  unsigned long long* keys, int* vals) {
    const long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const long long i = s.lo + q;
    if (i >= s.hi) return;
    const double x[3] = {s.x[0][i], s.x[1][i], s.x[2][i]};
    const double v[3] = {s.v[0][i], s.v[1][i], s.v[2][i]};
    unsigned long long key = ~0ull;
    if (body_live(x, v, s.mass[i])) {
        constexpr int kBits = DIMS == 3 ? 21 : 31;
        const double half = (double)(1u << (kBits - 1));
        const double inv = pow2d(kBits - 1 - extent_exponent(extent[slot]));  // |x| < 2^ex -> |x * inv| < 2^(bits-1)
        unsigned int c[3] = {0u, 0u, 0u};
#pragma unroll
        for (int k = 0; k < DIMS; ++k)
            c[k] = (unsigned int)fmin(fmax(__dmul_rn(x[k], inv) + half, 0.0), 2.0 * half - 1.0);
        key = hilbert_key<DIMS>(c, kBits);
    }
    keys[q] = key;
    vals[q] = (int)i;
}

__global__ void __launch_bounds__(kStateThreads) iota_order_kernel(int* order, long long lo, long long hi) {
    const long long i = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i < hi) order[i] = (int)i;
}

// FP64 state -> group-relative FP32 j-stream (hi, lo, group headers) + the gather filter's flags,
// one CTA per group of kGroupBodies stream rows in [lo, row_cap). DIAG: the stream of the stand-alone
// compute_diagnostics (Physics.hpp:117-161): every body with a finite position takes part with its raw mass.
struct TxOut {
    double* x[3];
};
template <bool DIAG>
__global__ void __launch_bounds__(kGroupBodies) pack_kernel(StateView s, const int* __restrict__ order, PeerSet out,
                                                            TxOut tx, uint8_t* live_out, const unsigned int* extent,
                                                            int slot, int mass_exp, double g, float softening) {
    __shared__ float sbox[kGroupBodies / 32][6];
    const Scale sc = scale_for(extent[slot], mass_exp, g, softening);
    const double inv_mass = pow2d(-mass_exp);
    const long long row = s.lo + blockIdx.x * (long long)kGroupBodies + threadIdx.x;
    double x[3] = {0.0, 0.0, 0.0};
    float m = 0.f;
    bool attracts = false;
    if (row < s.hi) {
        const long long i = order[row];
        const double v[3] = {s.v[0][i], s.v[1][i], s.v[2][i]};
#pragma unroll
        for (int k = 0; k < 3; ++k) x[k] = s.x[k][i];
        m = s.mass[i];
        if (DIAG) {
            attracts = finite_d(x[0]) && finite_d(x[1]) && finite_d(x[2]);
        } else {
            attracts = body_live(x, v, m);
            live_out[i] = attracts ? 1 : 0;
        }
    }
    pack_group<kGroupBodies>(out, tx.x, row, x, m, attracts, sc.inv_len, inv_mass, sbox);
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
