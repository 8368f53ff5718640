// exact.cuh — FP64 "exact" force path: the reference's arithmetic, operation for operation.
//
// Physics::compute_gravity's direct branch (src/systems/Physics.hpp:231-258) adds, for body t,
// its partners in ascending index; the pair (i<j) gives body i  +((G*m_j)*d)*invR3 and body j
// -((G*m_i)*d)*invR3 with d = x_j - x_i. Negation and the swapped difference are exact in
// IEEE arithmetic, so BOTH cases equal ((G*m_k)*(x_k - x_t))*invR3 and one target-major loop
// with non-contracted FP64 mul/add, IEEE sqrt and division reproduces the reference's
// accelerations bit for bit. Used when bit-level parity matters more than throughput: small
// worlds (the three-body seed is dynamically unstable — any rounding difference is amplified
// e-fold every ~600 steps) and as the on-device FP64 checker of the FP32 path.
#pragma once
#include "epilogue.cuh"
#include "state.cuh"

namespace nb {

constexpr int kExactThreads = 128;

__global__ void __launch_bounds__(kExactThreads) exact_force_kernel(StateView s, double g, float softening) {
    __shared__ double sx[kExactThreads], sy[kExactThreads], sz[kExactThreads], sgm[kExactThreads];
    __shared__ unsigned char slive[kExactThreads];
    const double eps2 = __dmul_rn((double)softening, (double)softening);  // Physics.hpp:186
    const long long t = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const bool valid = t < s.hi;
    double xt[3] = {0.0, 0.0, 0.0}, vt[3] = {0.0, 0.0, 0.0};
    float mt = 0.f;
    if (valid) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            xt[k] = s.x[k][t];
            vt[k] = s.v[k][t];
        }
        mt = s.mass[t];
    }
    const bool live_t = valid && body_live(xt, vt, mt);
    double ax = 0.0, ay = 0.0, az = 0.0;
    for (long long j0 = 0; j0 < s.n; j0 += kExactThreads) {
        const long long j = j0 + threadIdx.x;
        if (j < s.n) {
            const double xj[3] = {s.x[0][j], s.x[1][j], s.x[2][j]};
            const double vj[3] = {s.v[0][j], s.v[1][j], s.v[2][j]};
            const float mj = s.mass[j];
            sx[threadIdx.x] = xj[0];
            sy[threadIdx.x] = xj[1];
            sz[threadIdx.x] = xj[2];
            sgm[threadIdx.x] = __dmul_rn(g, (double)mj);  // G * m_j, :243
            slive[threadIdx.x] = body_live(xj, vj, mj) ? 1 : 0;
        } else {
            slive[threadIdx.x] = 0;
        }
        __syncthreads();
        if (live_t) {
            const int cnt = (int)min((long long)kExactThreads, s.n - j0);
            for (int u = 0; u < cnt; ++u) {
                if (!slive[u] || j0 + u == t) continue;
                const double dx = __dsub_rn(sx[u], xt[0]);
                const double dy = __dsub_rn(sy[u], xt[1]);
                const double dz = __dsub_rn(sz[u], xt[2]);
                double r2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
                r2 = __dadd_rn(r2, __dmul_rn(dz, dz));  // +0.0 in 2-D: exact
                r2 = __dadd_rn(r2, eps2);               // :239
                const double inv_r = __ddiv_rn(1.0, __dsqrt_rn(r2));            // :240
                const double inv_r3 = __dmul_rn(__dmul_rn(inv_r, inv_r), inv_r);  // :241
                const double gm = sgm[u];
                ax = __dadd_rn(ax, __dmul_rn(__dmul_rn(gm, dx), inv_r3));
                ay = __dadd_rn(ay, __dmul_rn(__dmul_rn(gm, dy), inv_r3));
                az = __dadd_rn(az, __dmul_rn(__dmul_rn(gm, dz), inv_r3));
            }
        }
        __syncthreads();
    }
    if (live_t) {
        const bool pinned = s.pinned[t] != 0;  // :248-255: pinned bodies keep a = 0
        s.a[0][t] = pinned ? 0.0 : ax;
        s.a[1][t] = pinned ? 0.0 : ay;
        s.a[2][t] = pinned ? 0.0 : az;
        s.live[t] = 1;
    } else if (valid) {
        s.live[t] = 0;
    }
}

// Velocity Verlet kick, Physics.hpp:309-323.
__global__ void __launch_bounds__(kStateThreads) verlet_kick_kernel(StateView s, double dt, float max_speed) {
    const long long i = s.lo + blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= s.hi || s.pinned[i]) return;
    double v[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double am = __dmul_rn(__dadd_rn(s.a0[k][i], s.a[k][i]), 0.5);
        v[k] = __dadd_rn(s.v[k][i], __dmul_rn(am, dt));
    }
    cap_speed(v, max_speed);
#pragma unroll
    for (int k = 0; k < 3; ++k) s.v[k][i] = v[k];
}

}  // namespace nb
