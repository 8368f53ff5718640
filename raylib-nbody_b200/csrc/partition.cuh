// partition.cuh — which rank owns which body (multi-GPU contexts).
//
// The C-ABI partitions a world by UPLOAD order: rank r uploads and downloads rows [lo, hi) of the caller's arrays
// (nb_partition). Owning the same rows as targets would give every rank a random 1/P sample of the world when the
// caller's order is not spatial (Plummer generators, the reference's entity order): its 128-row groups are then
// P^(1/3) times larger than a single GPU's and several times more (warp, group) visits take the two-float near
// path. So ownership is decoupled from the interface: every rank holds the same permutation `perm` (internal row
// -> upload row) that lists all bodies along ONE Hilbert curve of the uploaded positions, and rank r owns internal
// rows [lo, hi) — a compact stretch of the curve. Upload-order rows travel between ranks through an exchange buffer
// XB (12 SoA arrays x, v, a, a0 + mass + pinned flags, all rows on every rank): filled by NCCL all-gathers
// (one process per GPU) or by kernels that store to every member's copy over NVLink (one process, several GPUs),
// then read with `perm` (upload) or `inv_perm` (download). A single-GPU context never comes here.
#pragma once
#include "state.cuh"

namespace nb {

constexpr int kPartitionExtentSlot = kExtentSlots + 1;
constexpr int kXbArrays = 12;  // x[3], v[3], a[3], a0[3]

struct XbOut {  // every device's exchange buffer (own first), for kernels that publish rows to all of them
    int n;
    double* st[8];
    float* mass[8];
    uint8_t* pinned[8];
};

// staging AoS [pos dim*rows][vel][acc] of upload rows [lo, lo + rows) -> SoA arrays 0..8 of every XB in `out`;
// mass / pinned of those rows are already in out.mass[0] / out.pinned[0] (copied from the host) and are forwarded.
__global__ void __launch_bounds__(kStateThreads) unpack_aos_xb_kernel(const double* pos, const double* vel,
                                                                      const double* acc, int dim, long long lo,
                                                                      long long rows, long long cap, XbOut out) {
    const long long q = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (q >= rows) return;
    const long long i = lo + q;
    double w[9];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        w[k] = k < dim ? pos[q * dim + k] : 0.0;
        w[3 + k] = k < dim ? vel[q * dim + k] : 0.0;
        w[6 + k] = (acc && k < dim) ? acc[q * dim + k] : 0.0;
    }
    const float m = out.mass[0][i];
    const uint8_t pin = out.pinned[0][i];
    for (int t = 0; t < out.n; ++t) {
#pragma unroll
        for (int a = 0; a < 9; ++a) out.st[t][(long long)a * cap + i] = w[a];
        if (t > 0) { out.mass[t][i] = m; out.pinned[t][i] = pin; }
    }
}

// a member's own internal rows [lo, hi) of the first `n_arrays` state arrays (+ mass, pinned) -> every XB
__global__ void __launch_bounds__(kStateThreads) publish_state_kernel(StateView s, int n_arrays, bool with_mass,
                                                                      long long cap, XbOut out) {
    const long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= s.hi) return;
    for (int a = 0; a < n_arrays; ++a) {
        double* const* src = a < 3 ? s.x : (a < 6 ? s.v : (a < 9 ? s.a : s.a0));
        const double w = src[a % 3][i];
        for (int t = 0; t < out.n; ++t) out.st[t][(long long)a * cap + i] = w;
    }
    if (with_mass) {
        const float m = s.mass[i];
        const uint8_t pin = s.pinned[i];
        for (int t = 0; t < out.n; ++t) { out.mass[t][i] = m; out.pinned[t][i] = pin; }
    }
}

// max |coordinate| over the bodies of XB that pass the gather filter -> extent[kPartitionExtentSlot]
__global__ void __launch_bounds__(kStateThreads) xb_extent_kernel(const double* xb, long long cap, long long n,
                                                                  const float* mass, unsigned int* extent) {
    __shared__ float red[kStateThreads / 32];
    float e = 0.f;
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double x[3] = {xb[i], xb[cap + i], xb[2 * cap + i]};
        const double v[3] = {xb[3 * cap + i], xb[4 * cap + i], xb[5 * cap + i]};
        e = fmaxf(e, extent_of(x, body_live(x, v, mass[i])));
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) e = fmaxf(e, __shfl_xor_sync(0xffffffffu, e, o));
    if ((threadIdx.x & 31) == 0) red[threadIdx.x >> 5] = e;
    __syncthreads();
    if (threadIdx.x == 0) {
        for (int w = 1; w < kStateThreads / 32; ++w) e = fmaxf(e, red[w]);
        atomicMax(&extent[kPartitionExtentSlot], __float_as_uint(e));
    }
}

// Hilbert key of every body, ~0 for bodies the gather filter drops: they sort to the end of the curve and land on
// the last rank(s). Pair j is the body of UPLOAD row j, whatever order XB is in (row_of == nullptr: XB is in upload
// order; else XB row row_of[j]), and vals[j] = its XB row: the stable sort then breaks ties between bodies of one
// curve cell by upload row, so the permutation depends on the positions only, not on who held the rows before
// (a resumed run re-partitions exactly like the uninterrupted one).
template <int DIMS>
__global__ void __launch_bounds__(kStateThreads) xb_curve_key_kernel(const double* xb, long long cap, long long n,
                                                                     const float* mass, const int* row_of,
                                                                     const unsigned int* extent,
                                                                     unsigned long long* keys, int* vals) {
    const long long j = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (j >= n) return;
    const long long i = row_of ? row_of[j] : j;
    const double x[3] = {xb[i], xb[cap + i], xb[2 * cap + i]};
    const double v[3] = {xb[3 * cap + i], xb[4 * cap + i], xb[5 * cap + i]};
    unsigned long long key = ~0ull;
    if (body_live(x, v, mass[i])) {
        constexpr int kBits = DIMS == 3 ? 21 : 31;
        const double half = (double)(1u << (kBits - 1));
        const double inv = pow2d(kBits - 1 - extent_exponent(extent[kPartitionExtentSlot]));
        unsigned int c[3] = {0u, 0u, 0u};
#pragma unroll
        for (int k = 0; k < DIMS; ++k)
            c[k] = (unsigned int)fmin(fmax(__dmul_rn(x[k], inv) + half, 0.0), 2.0 * half - 1.0);
        key = hilbert_key<DIMS>(c, kBits);
    }
    keys[j] = key;
    vals[j] = (int)i;
}

// perm[k] = through ? through[sorted[k]] : sorted[k];  inv[perm[k]] = k
__global__ void __launch_bounds__(kStateThreads) set_perm_kernel(const int* sorted, const int* through, long long n,
                                                                 int* perm, int* inv) {
    const long long k = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (k >= n) return;
    const int b = through ? through[sorted[k]] : sorted[k];
    perm[k] = b;
    inv[b] = (int)k;
}

// own internal rows k in [lo, hi): state arrays 0..n_arrays-1 (the rest zero), mass and pinned from XB row
// via[perm[k]] (via == nullptr: XB is in upload order; via == the previous inv_perm: XB is in the previous
// internal order)
__global__ void __launch_bounds__(kStateThreads) gather_owned_kernel(const double* xb, const float* xb_mass,
                                                                     const uint8_t* xb_pinned, long long cap,
                                                                     const int* perm, const int* via, int n_arrays,
                                                                     StateView s, float* mass, uint8_t* pinned) {
    const long long k = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (k >= s.hi) return;
    long long b = perm[k];
    if (via) b = via[b];
#pragma unroll
    for (int a = 0; a < kXbArrays; ++a) {
        double* const* dst = a < 3 ? s.x : (a < 6 ? s.v : (a < 9 ? s.a : s.a0));
        dst[a % 3][k] = a < n_arrays ? xb[(long long)a * cap + b] : 0.0;
    }
    mass[k] = xb_mass[b];
    pinned[k] = xb_pinned[b];
}

// XB holds array group `which` (0 x, 1 v, 2 a, 3 a0) of every body in INTERNAL order -> AoS rows of upload rows
// [lo, hi): out[(b - lo) * dim + k] = XB[which * 3 + k][inv_perm[b]]
__global__ void __launch_bounds__(kStateThreads) pack_aos_xb_kernel(const double* xb, long long cap, const int* inv_perm,
                                                                    int which, int dim, long long lo, long long hi,
                                                                    double* out) {
    const long long b = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (b >= hi) return;
    const long long k = inv_perm[b];
    for (int c = 0; c < dim; ++c) out[(b - lo) * dim + c] = xb[(long long)(which * 3 + c) * cap + k];
}

// stream order in upload-row terms and back: out[q] = map[in[q]]
__global__ void __launch_bounds__(kStateThreads) map_rows_kernel(const int* in, const int* map, long long lo,
                                                                 long long hi, int* out) {
    const long long q = lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (q >= hi) return;
    out[q] = map[in[q]];
}

}  // namespace nb
