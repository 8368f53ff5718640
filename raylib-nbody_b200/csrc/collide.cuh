// collide.cuh — the pair tests of the reference's collision pass on the device.
//
// Collision::resolve (src/systems/Collision.hpp:49-203) snapshots every body that passes its
// input filter (finite position, finite mass > 0, :63-65) and walks all pairs i < j testing
//     dx = xj - xi; dy = yj - yi; rsum = ri + rj; dist2 = dx*dx + dy*dy;
//     if (dist2 > rsum * rsum) continue;            // no overlap            (:76-84)
// merging the pairs that overlap as it goes. When no pair overlaps the whole pass is a no-op — the
// common frame, and an O(N^2) FP64 loop on one host core (22 ms at N = 4096, ~25 min at 2^20). The
// kernels here evaluate exactly that predicate, in the reference's operation order with
// non-contracted FP64 arithmetic: over all pairs (count, first pair, optionally the list of pairs:
// nb_count_overlaps and the first step of nb_resolve_collisions); the merges themselves are replayed
// in the reference's order by merge_replay_kernel (merge.cuh), which re-tests pairs with this predicate.
//
// FP64-pipe bound: 8 FP64 operations per pair; triangular sweep (j-tiles at or after the
// target's own tile).
#pragma once
#include "state.cuh"

namespace nb {

constexpr int kOverlapThreads = 256;

__device__ __forceinline__ bool collide_input_ok(double x, double y, double z, float m) {
    return finite_d(x) && finite_d(y) && finite_d(z) && m > 0.0f && m <= 3.4028234663852886e38f;
}

// result[0] = number of overlapping pairs, result[1] = min over them of i * n + j. With a pair sink
// (nb_resolve_collisions) every overlapping pair's key i * n + j is also appended to pairs[] (result[2]
// counts the appends; keys beyond pair_cap are dropped and the caller retries with a larger buffer).
__device__ __forceinline__ void emit_pair(unsigned long long key, unsigned long long* pairs, unsigned long long pair_cap,
                                          unsigned long long* result) {
    if (!pairs) return;
    const unsigned long long slot = atomicAdd(&result[2], 1ull);
    if (slot < pair_cap) pairs[slot] = key;
}

template <bool D2>
__global__ void __launch_bounds__(kOverlapThreads) overlap_kernel(const double* __restrict__ x0,
                                                                  const double* __restrict__ x1,
                                                                  const double* __restrict__ x2,
                                                                  const float* __restrict__ mass,
                                                                  const double* __restrict__ radius, long long n,
                                                                  unsigned long long* result,
                                                                  unsigned long long* pairs = nullptr,
                                                                  unsigned long long pair_cap = 0) {
    __shared__ double sx[kOverlapThreads], sy[kOverlapThreads], sz[kOverlapThreads], sr[kOverlapThreads];
    __shared__ unsigned char sok[kOverlapThreads];
    const long long i = blockIdx.x * (long long)kOverlapThreads + threadIdx.x;
    double xi = 0.0, yi = 0.0, zi = 0.0, ri = 0.0;
    bool ok_i = false;
    if (i < n) {
        xi = x0[i]; yi = x1[i]; zi = x2[i]; ri = radius[i];
        ok_i = collide_input_ok(xi, yi, zi, mass[i]);
    }
    unsigned long long count = 0, first = ~0ull;
    for (long long j0 = blockIdx.x * (long long)kOverlapThreads; j0 < n; j0 += kOverlapThreads) {
        const long long jl = j0 + threadIdx.x;
        if (jl < n) {
            const double xj = x0[jl], yj = x1[jl], zj = x2[jl];
            const bool ok = collide_input_ok(xj, yj, zj, mass[jl]);
            // filtered sources sit at +inf: dist2 = +inf > rsum^2 for every finite radius, so the
            // hot loop needs no flag test; sok[] settles the (rare) candidates exactly
            sx[threadIdx.x] = ok ? xj : __longlong_as_double(0x7ff0000000000000ll);
            sy[threadIdx.x] = ok ? yj : 0.0;
            sz[threadIdx.x] = ok ? zj : 0.0;
            sr[threadIdx.x] = radius[jl];
            sok[threadIdx.x] = ok ? 1 : 0;
        } else {
            sx[threadIdx.x] = __longlong_as_double(0x7ff0000000000000ll);
            sy[threadIdx.x] = sz[threadIdx.x] = sr[threadIdx.x] = 0.0;
            sok[threadIdx.x] = 0;
        }
        __syncthreads();
        if (ok_i) {
#pragma unroll 8
            for (int u = 0; u < kOverlapThreads; ++u) {
                const double dx = __dsub_rn(sx[u], xi);
                const double dy = __dsub_rn(sy[u], yi);
                const double rsum = __dadd_rn(ri, sr[u]);
                double dist2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
                if (!D2) {  // reference (2-D) uploads have z == 0: the term would add +0.0 exactly
                    const double dz = __dsub_rn(sz[u], zi);
                    dist2 = __dadd_rn(dist2, __dmul_rn(dz, dz));
                }
                if (!(dist2 > __dmul_rn(rsum, rsum))) {
                    const long long j = j0 + u;
                    if (j > i && j < n && sok[u]) {
                        ++count;
                        const unsigned long long key = (unsigned long long)i * (unsigned long long)n + (unsigned long long)j;
                        first = key < first ? key : first;
                        emit_pair(key, pairs, pair_cap, result);
                    }
                }
            }
        }
        __syncthreads();
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        count += __shfl_xor_sync(0xffffffffu, count, o);
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, first, o);
        first = other < first ? other : first;
    }
    if ((threadIdx.x & 31) == 0 && count) {
        atomicAdd(&result[0], count);
        atomicMin(&result[1], first);
    }
}

// ---- O(N) broad phase: uniform grid with cells of 2 r_max -----------------------------------------
// Two bodies can only overlap if their distance is <= r_i + r_j <= 2 r_max, so with cells of edge
// h = 2 r_max (1 + 1e-9) an overlapping pair always sits in the same or in adjacent cells (the margin
// covers the rounding of the cell coordinate for up to 2^20 cells per axis and of dist2 itself). Bodies
// are binned by a counting sort (atomic histogram, exclusive scan, atomic fill) and each body tests the
// EXACT predicate of overlap_kernel against the bodies of its 3^dim neighbourhood with a larger index.
// Counts and the minimum (i, j) key are order-independent, so the result equals the all-pairs sweep's.
struct GridSpec {
    double origin[3];
    double inv_h;
    int cells[3];   // per axis (1 on unused axes)
    int dim;
};

__device__ __forceinline__ unsigned long long dbl_order_key(double v) {  // monotone map double -> u64
    const unsigned long long b = (unsigned long long)__double_as_longlong(v);
    return (b >> 63) ? ~b : (b | 0x8000000000000000ull);
}
__host__ __device__ inline double dbl_from_order_key(unsigned long long k) {
    const unsigned long long b = (k >> 63) ? (k & 0x7fffffffffffffffull) : ~k;
    union { unsigned long long u; double d; } c;
    c.u = b;
    return c.d;
}

// bounds[2k] = min key, bounds[2k+1] = max key of coordinate k over bodies passing the input filter
__global__ void __launch_bounds__(256) collide_bounds_kernel(const double* __restrict__ x0, const double* __restrict__ x1,
                                                             const double* __restrict__ x2, const float* __restrict__ mass,
                                                             long long n, unsigned long long* bounds) {
    unsigned long long lo[3] = {~0ull, ~0ull, ~0ull}, hi[3] = {0ull, 0ull, 0ull};
    for (long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x; i < n; i += (long long)gridDim.x * blockDim.x) {
        const double p[3] = {x0[i], x1[i], x2[i]};
        if (!collide_input_ok(p[0], p[1], p[2], mass[i])) continue;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            const unsigned long long key = dbl_order_key(p[k]);
            lo[k] = key < lo[k] ? key : lo[k];
            hi[k] = key > hi[k] ? key : hi[k];
        }
    }
#pragma unroll
    for (int k = 0; k < 3; ++k) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            const unsigned long long a = __shfl_xor_sync(0xffffffffu, lo[k], o), b = __shfl_xor_sync(0xffffffffu, hi[k], o);
            lo[k] = a < lo[k] ? a : lo[k];
            hi[k] = b > hi[k] ? b : hi[k];
        }
        if ((threadIdx.x & 31) == 0) {
            atomicMin(&bounds[2 * k], lo[k]);
            atomicMax(&bounds[2 * k + 1], hi[k]);
        }
    }
}

__device__ __forceinline__ int grid_axis_cell(double x, double origin, double inv_h, int cells) {
    const double c = floor(__dmul_rn(__dsub_rn(x, origin), inv_h));
    const int q = c < 0.0 ? 0 : (c >= (double)cells ? cells - 1 : (int)c);
    return q;
}

__device__ __forceinline__ long long grid_cell_of(const GridSpec& g, double x, double y, double z) {
    const int cx = grid_axis_cell(x, g.origin[0], g.inv_h, g.cells[0]);
    const int cy = grid_axis_cell(y, g.origin[1], g.inv_h, g.cells[1]);
    const int cz = g.dim == 3 ? grid_axis_cell(z, g.origin[2], g.inv_h, g.cells[2]) : 0;
    return ((long long)cz * g.cells[1] + cy) * g.cells[0] + cx;
}

// pass 1 (fill == nullptr): histogram; pass 2: scatter body indices into their cell's slice
__global__ void __launch_bounds__(256) grid_bin_kernel(const double* __restrict__ x0, const double* __restrict__ x1,
                                                       const double* __restrict__ x2, const float* __restrict__ mass,
                                                       long long n, GridSpec g, unsigned int* cell_count,
                                                       const unsigned int* cell_start, unsigned int* cell_fill,
                                                       unsigned int* sorted_idx) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    const double x = x0[i], y = x1[i], z = x2[i];
    if (!collide_input_ok(x, y, z, mass[i])) return;
    const long long c = grid_cell_of(g, x, y, z);
    if (!cell_fill) {
        atomicAdd(&cell_count[c], 1u);
    } else {
        const unsigned int slot = atomicAdd(&cell_fill[c], 1u);
        sorted_idx[cell_start[c] + slot] = (unsigned int)i;
    }
}

__global__ void __launch_bounds__(256) grid_overlap_kernel(const double* __restrict__ x0, const double* __restrict__ x1,
                                                           const double* __restrict__ x2, const float* __restrict__ mass,
                                                           const double* __restrict__ radius, long long n, GridSpec g,
                                                           const unsigned int* __restrict__ cell_start,
                                                           const unsigned int* __restrict__ sorted_idx,
                                                           unsigned long long* result,
                                                           unsigned long long* pairs = nullptr,
                                                           unsigned long long pair_cap = 0) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    unsigned long long count = 0, first = ~0ull;
    if (i < n) {
        const double xi = x0[i], yi = x1[i], zi = x2[i], ri = radius[i];
        if (collide_input_ok(xi, yi, zi, mass[i])) {
            const int cx = grid_axis_cell(xi, g.origin[0], g.inv_h, g.cells[0]);
            const int cy = grid_axis_cell(yi, g.origin[1], g.inv_h, g.cells[1]);
            const int cz = g.dim == 3 ? grid_axis_cell(zi, g.origin[2], g.inv_h, g.cells[2]) : 0;
            const int z_lo = g.dim == 3 ? max(cz - 1, 0) : 0, z_hi = g.dim == 3 ? min(cz + 1, g.cells[2] - 1) : 0;
            for (int qz = z_lo; qz <= z_hi; ++qz)
                for (int qy = max(cy - 1, 0); qy <= min(cy + 1, g.cells[1] - 1); ++qy) {
                    // the row's up-to-three cells are contiguous in the sorted array
                    const long long row = ((long long)qz * g.cells[1] + qy) * g.cells[0];
                    const unsigned int k_lo = cell_start[row + max(cx - 1, 0)];
                    const unsigned int k_hi = cell_start[row + min(cx + 1, g.cells[0] - 1) + 1];
                    for (unsigned int k = k_lo; k < k_hi; ++k) {
                        const long long j = sorted_idx[k];
                        if (j <= i) continue;  // each pair once, as (i < j)
                        const double dx = __dsub_rn(x0[j], xi);
                        const double dy = __dsub_rn(x1[j], yi);
                        const double rsum = __dadd_rn(ri, radius[j]);
                        double dist2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
                        if (g.dim == 3) {
                            const double dz = __dsub_rn(x2[j], zi);
                            dist2 = __dadd_rn(dist2, __dmul_rn(dz, dz));
                        }
                        if (!(dist2 > __dmul_rn(rsum, rsum))) {
                            ++count;
                            const unsigned long long key = (unsigned long long)i * (unsigned long long)n + (unsigned long long)j;
                            first = key < first ? key : first;
                            emit_pair(key, pairs, pair_cap, result);
                        }
                    }
                }
        }
    }
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) {
        count += __shfl_xor_sync(0xffffffffu, count, o);
        const unsigned long long other = __shfl_xor_sync(0xffffffffu, first, o);
        first = other < first ? other : first;
    }
    if ((threadIdx.x & 31) == 0 && count) {
        atomicAdd(&result[0], count);
        atomicMin(&result[1], first);
    }
}

}  // namespace nb
