// collide.cuh — detection phase of the reference's collision pass on the device.
//
// Collision::resolve (src/systems/Collision.hpp:49-203) snapshots every body that passes its
// input filter (finite position, finite mass > 0, :63-65) and walks all pairs i < j testing
//     dx = xj - xi; dy = yj - yi; rsum = ri + rj; dist2 = dx*dx + dy*dy;
//     if (dist2 > rsum * rsum) continue;            // no overlap            (:76-84)
// before it merges anything. When no pair overlaps the whole pass is a no-op — the common frame,
// and an O(N^2) FP64 loop on one host core (22 ms at N = 4096, ~25 min at 2^20). This kernel
// evaluates exactly that predicate, in the reference's operation order with non-contracted FP64
// arithmetic, for all pairs i < j, and returns the number of overlapping pairs and the first one
// in the reference's visiting order. The caller (include/reference_side/PhysicsB200.hpp) runs the
// reference's own resolve only on frames where the count is non-zero, so merge semantics stay
// the reference's, bit for bit.
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

// result[0] = number of overlapping pairs, result[1] = min over them of i * n + j
template <bool D2>
__global__ void __launch_bounds__(kOverlapThreads) overlap_kernel(const double* __restrict__ x0,
                                                                  const double* __restrict__ x1,
                                                                  const double* __restrict__ x2,
                                                                  const float* __restrict__ mass,
                                                                  const double* __restrict__ radius, long long n,
                                                                  unsigned long long* result) {
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

}  // namespace nb
