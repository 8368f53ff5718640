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

// Partners [0,cnt) of a shared-memory tile added to (ax,ay,az) in ascending index, skipping
// filtered partners and the target itself (self = its index in the tile, or -1). The pair terms of
// kExactIlp consecutive partners are evaluated independently (the sqrt/div chains overlap) and then
// added strictly in order, each add predicated on the partner's flag: same operations, same order
// and the same skipped adds as a one-by-one loop, so the sums are bit-identical to it.
constexpr int kExactIlp = 1;  // measured on B200: 8 does not help (n = 256 is FP64-throughput-bound on one SM) and slows n = 3

__device__ __forceinline__ void exact_accumulate(const double* sx, const double* sy, const double* sz, const double* sgm,
                                                 const unsigned char* slive, int cnt, int self, double x0, double x1,
                                                 double x2, double eps2, double& ax, double& ay, double& az) {
    for (int u0 = 0; u0 < cnt; u0 += kExactIlp) {
        double tx[kExactIlp], ty[kExactIlp], tz[kExactIlp];
        bool use[kExactIlp];
#pragma unroll
        for (int q = 0; q < kExactIlp; ++q) {
            const int u = u0 + q;
            use[q] = u < cnt && slive[u] && u != self;
            const double dx = __dsub_rn(sx[u], x0);
            const double dy = __dsub_rn(sy[u], x1);
            const double dz = __dsub_rn(sz[u], x2);
            double r2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
            r2 = __dadd_rn(r2, __dmul_rn(dz, dz));  // +0.0 in 2-D: exact
            r2 = __dadd_rn(r2, eps2);               // Physics.hpp:239
            const double inv_r = __ddiv_rn(1.0, __dsqrt_rn(r2));            // :240
            const double inv_r3 = __dmul_rn(__dmul_rn(inv_r, inv_r), inv_r);  // :241
            const double gm = sgm[u];
            tx[q] = __dmul_rn(__dmul_rn(gm, dx), inv_r3);
            ty[q] = __dmul_rn(__dmul_rn(gm, dy), inv_r3);
            tz[q] = __dmul_rn(__dmul_rn(gm, dz), inv_r3);
        }
#pragma unroll
        for (int q = 0; q < kExactIlp; ++q) {
            if (use[q]) {
                ax = __dadd_rn(ax, tx[q]);
                ay = __dadd_rn(ay, ty[q]);
                az = __dadd_rn(az, tz[q]);
            }
        }
    }
}

__global__ void __launch_bounds__(kExactThreads) exact_force_kernel(StateView s, double g, float softening) {
    __shared__ double sx[kExactThreads], sy[kExactThreads], sz[kExactThreads], sgm[kExactThreads];
    __shared__ unsigned char slive[kExactThreads];
    static_assert(kExactThreads % kExactIlp == 0, "tile reads stay inside the arrays");
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
            const long long self = t - j0;
            exact_accumulate(sx, sy, sz, sgm, slive, cnt, self >= 0 && self < cnt ? (int)self : -1, xt[0], xt[1], xt[2],
                             eps2, ax, ay, az);
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

// ---- small worlds: every substep of a call in ONE launch -----------------------------------------
// The reference's interactive sizes (3 ... a few hundred bodies) are launch-bound on a GPU: three
// kernels per Verlet substep at ~4 us each against ~1 us of arithmetic. For n <= kSmallWorldMax a
// single CTA owns the whole world — one thread per body with its state in registers, positions
// mirrored in shared memory — and loops over all substeps with CTA barriers in between. The
// arithmetic is the same non-contracted FP64 sequence as exact_force_kernel / verlet_drift_kernel /
// verlet_kick_kernel / euler_move_kernel, so results are bit-identical to the multi-launch path
// (and to the reference).
constexpr int kSmallWorldMax = 512;
static_assert(kSmallWorldMax % kExactIlp == 0, "tile reads stay inside the arrays");

// first_force: evaluate a(x) before the first substep (Verlet with Acceleration != a(x); always for Euler,
// whose frame is gravity -> move). integrator: 0 Euler (force, move) x n_sub; 1 Verlet (drift, force, kick) x n_sub.
__global__ void __launch_bounds__(kSmallWorldMax) small_world_steps_kernel(StateView s, double g, float softening,
                                                                           double dt, double half_dt2, float max_speed,
                                                                           int integrator, int first_force,
                                                                           long long n_sub) {
    __shared__ double sx[kSmallWorldMax], sy[kSmallWorldMax], sz[kSmallWorldMax], sgm[kSmallWorldMax];
    __shared__ unsigned char slive[kSmallWorldMax];
    const int t = threadIdx.x;
    const int n = (int)s.n;
    const bool valid = t < n;
    const double eps2 = __dmul_rn((double)softening, (double)softening);  // Physics.hpp:186
    double x[3] = {0.0, 0.0, 0.0}, v[3] = {0.0, 0.0, 0.0}, a[3] = {0.0, 0.0, 0.0}, a0[3] = {0.0, 0.0, 0.0};
    float m = 0.f;
    bool pinned = false;
    if (valid) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            x[k] = s.x[k][t]; v[k] = s.v[k][t]; a[k] = s.a[k][t]; a0[k] = s.a0[k][t];
        }
        m = s.mass[t];
        pinned = s.pinned[t] != 0;
        sgm[t] = __dmul_rn(g, (double)m);  // G * m_j, Physics.hpp:243
    }
    bool live = false;

    // publish this body's position and gather-filter flag, then a(x) over all partners in ascending index
    auto force = [&]() {
        live = valid && body_live(x, v, m);
        if (valid) {
            sx[t] = x[0]; sy[t] = x[1]; sz[t] = x[2];
            slive[t] = live ? 1 : 0;
        }
        __syncthreads();
        if (live) {
            double ax = 0.0, ay = 0.0, az = 0.0;
            exact_accumulate(sx, sy, sz, sgm, slive, n, t, x[0], x[1], x[2], eps2, ax, ay, az);
            a[0] = pinned ? 0.0 : ax;  // Physics.hpp:248-255
            a[1] = pinned ? 0.0 : ay;
            a[2] = pinned ? 0.0 : az;
        }
        __syncthreads();  // everyone is done reading before the next move overwrites the mirror
    };

    if (first_force && integrator != 0) force();
    for (long long k = 0; k < n_sub; ++k) {
        if (integrator == 0) {
            force();  // the gravity system (or the refresh of Physics.hpp:292), then the move of :276-290
            if (valid && !pinned) {
#pragma unroll
                for (int c = 0; c < 3; ++c) v[c] = __dadd_rn(v[c], __dmul_rn(a[c], dt));
                cap_speed(v, max_speed);
#pragma unroll
                for (int c = 0; c < 3; ++c) x[c] = __dadd_rn(x[c], __dmul_rn(v[c], dt));
            }
        } else {
            if (valid && !pinned) {  // drift, Physics.hpp:297-304
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    x[c] = __dadd_rn(x[c], __dadd_rn(__dmul_rn(v[c], dt), __dmul_rn(a[c], half_dt2)));
                    a0[c] = a[c];
                }
            }
            force();
            if (valid && !pinned) {  // kick, Physics.hpp:309-323
#pragma unroll
                for (int c = 0; c < 3; ++c) {
                    const double am = __dmul_rn(__dadd_rn(a0[c], a[c]), 0.5);
                    v[c] = __dadd_rn(v[c], __dmul_rn(am, dt));
                }
                cap_speed(v, max_speed);
            }
        }
    }
    if (valid) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            s.x[k][t] = x[k]; s.v[k][t] = v[k]; s.a[k][t] = a[k]; s.a0[k][t] = a0[k];
        }
        s.live[t] = body_live(x, v, m) ? 1 : 0;
    }
}

}  // namespace nb
