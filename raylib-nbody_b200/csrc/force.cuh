// force.cuh — the hot kernel: j-tiled direct-sum softened gravity with the integrator
// kick/drift (and the next step's FP32 pack) fused into its epilogue.
//
// Replaces the reference's hot loop A, src/systems/Physics.hpp:231-258 (direct all-pairs
// branch of compute_gravity) and the per-body loops of integrate(), Physics.hpp:276-290
// (semi-implicit Euler) and :297-323 (velocity Verlet). It is a full N x N target-major
// sweep (each target sums all sources in ascending index), not the reference's triangular
// pair loop, so no atomics and a fixed, reproducible summation order.
//
// Per interaction (3-D): 3 FADD, 3 FFMA (r^2 seeded with eps^2), 2 FMUL (r^6), 1 MUFU.RSQ
// (r^-3 directly: one rsqrt error instead of three), 1 FMUL (m r^-3), 3 FFMA  = 12 FP32-pipe
// ops + 1 SFU op = the conventional 20 FLOPs. With PACKED the 12 FP32 ops are issued as 6
// two-wide FFMA2/FADD2/FMUL2 on a pair of sources. Compute-bound on the FP32 pipe; HBM sees
// the j-stream once per CTA through L2 (TMA bulk copies into a shared-memory ring).
#pragma once
#include "common.cuh"
#include "ptx.cuh"

#ifndef NB_UNROLL_PAIRS
#define NB_UNROLL_PAIRS 32  // pair records per unrolled loop body (32 = a whole 64-source flush chunk)
#endif

namespace nb {

constexpr int kUnrollPairs = NB_UNROLL_PAIRS;

// One target body's running sums: FP32 inside a flush chunk, FP64 across chunks.
// EQM: all sources carry the same mass -> the m*w multiply is dropped (the common factor is
// applied in FP64 in the epilogue); D2: z == 0 for every body (reference 2-D uploads).
template <bool PACKED, bool POT, bool EQM = false, bool D2 = false>
struct Target;

// ---- scalar FP32 variant ------------------------------------------------------------
template <bool POT>
struct Target<false, POT, false, false> {
    float xi, yi, zi;
    float ax, ay, az, ph;
    __device__ __forceinline__ void init(float x, float y, float z) {
        xi = x; yi = y; zi = z;
        ax = ay = az = ph = 0.f;
    }
    // rec: one pair record {x0,x1,y0,y1 | z0,z1,m0,m1}; m0/m1 may be masked by the caller
    __device__ __forceinline__ void interact(const float4& A, const float4& B, float m0, float m1, float eps2) {
        one(A.x, A.z, B.x, m0, eps2);
        one(A.y, A.w, B.y, m1, eps2);
    }
    __device__ __forceinline__ void one(float xj, float yj, float zj, float mj, float eps2) {
        const float dx = xj - xi, dy = yj - yi, dz = zj - zi;
        float r2 = __fmaf_rn(dx, dx, eps2);
        r2 = __fmaf_rn(dy, dy, r2);
        r2 = __fmaf_rn(dz, dz, r2);
        const float r6 = (r2 * r2) * r2;
        const float w = rsqrt_fast(r6);  // r^-3
        const float s = mj * w;
        ax = __fmaf_rn(s, dx, ax);
        ay = __fmaf_rn(s, dy, ay);
        az = __fmaf_rn(s, dz, az);
        if (POT) ph = __fmaf_rn(mj, w * r2, ph);  // m / r
    }
    __device__ __forceinline__ void flush(double& Ax, double& Ay, double& Az, double& Ph) {
        Ax += (double)ax; Ay += (double)ay; Az += (double)az;
        ax = ay = az = 0.f;
        if (POT) { Ph += (double)ph; ph = 0.f; }
    }
};

// ---- packed f32x2 variant (two sources per instruction) -------------------------------
template <bool POT, bool EQM, bool D2>
struct Target<true, POT, EQM, D2> {
    f2 nxi, nyi, nzi;  // {-x,-x} ... so dx = xj + (-xi) is one FADD2
    f2 ax, ay, az, ph;
    __device__ __forceinline__ void init(float x, float y, float z) {
        nxi = f2_make(-x, -x); nyi = f2_make(-y, -y); nzi = f2_make(-z, -z);
        ax = ay = az = ph = f2_make(0.f, 0.f);
    }
    // mj: the sources' masses; for EQM kernels only read when MASKED, as a {0|1, 0|1} self-term mask
    template <bool MASKED = false>
    __device__ __forceinline__ void interact(f2 xj, f2 yj, f2 zj, f2 mj, f2 eps2) {
        const f2 dx = f2_add(xj, nxi), dy = f2_add(yj, nyi);
        f2 r2 = f2_fma(dx, dx, eps2);
        r2 = f2_fma(dy, dy, r2);
        f2 dz;
        if (!D2) {
            dz = f2_add(zj, nzi);
            r2 = f2_fma(dz, dz, r2);
        }
        const f2 r6 = f2_mul(f2_mul(r2, r2), r2);
        float lo, hi;
        f2_split(r6, lo, hi);
        const f2 w = f2_make(rsqrt_fast(lo), rsqrt_fast(hi));
        const f2 s = EQM ? w : f2_mul(mj, w);
        ax = f2_fma(s, dx, ax);
        ay = f2_fma(s, dy, ay);
        if (!D2) az = f2_fma(s, dz, az);
        if (POT) {
            if (EQM && !MASKED) ph = f2_fma(w, r2, ph);  // 1 / r (non-attracting rows: w = 0)
            else ph = f2_fma(mj, f2_mul(w, r2), ph);    // m / r
        }
    }
    // The even/odd-source halves are added in FP32 first: one commutative, sign-symmetric
    // rounding, so mirror-symmetric configurations (the reference's three-body seed,
    // Physics.hpp:87-115) keep exactly mirrored sums whatever slot a source landed in.
    __device__ __forceinline__ void flush(double& Ax, double& Ay, double& Az, double& Ph) {
        float lo, hi;
        f2_split(ax, lo, hi); Ax += (double)(lo + hi);
        f2_split(ay, lo, hi); Ay += (double)(lo + hi);
        if (!D2) { f2_split(az, lo, hi); Az += (double)(lo + hi); }
        ax = ay = az = f2_make(0.f, 0.f);
        if (POT) { f2_split(ph, lo, hi); Ph += (double)(lo + hi); ph = f2_make(0.f, 0.f); }
    }
};

// One flush chunk (CH sources = CH/2 pair records) against R targets.
// MASK: the chunk may contain one of this CTA's own targets; its mass is zeroed for that
// target so the potential carries no self term (the force self term is exactly 0 anyway).
template <int R, bool PACKED, bool POT, bool MASK, int CH, bool EQM = false, bool D2 = false>
__device__ __forceinline__ void chunk_interact(Target<PACKED, POT, EQM, D2> (&t)[R], const float4* __restrict__ rec,
                                               float eps2, long long j_first, const long long (&ti)[R]) {
    if constexpr (PACKED) {
        const f2 e2 = f2_make(eps2, eps2);
#pragma unroll kUnrollPairs
        for (int u = 0; u < CH / 2; ++u) {
            const ulonglong2 A = *reinterpret_cast<const ulonglong2*>(rec + 2 * u);
            f2 xj{A.x}, yj{A.y}, zj{0ull}, mj{0ull};
            constexpr bool kNeedM = !EQM;
            const unsigned long long* B64 = reinterpret_cast<const unsigned long long*>(rec + 2 * u + 1);
            if (!D2 && kNeedM) {
                const ulonglong2 B = *reinterpret_cast<const ulonglong2*>(rec + 2 * u + 1);
                zj.r = B.x;
                mj.r = B.y;
            } else if (!D2) {
                zj.r = B64[0];
            } else if (kNeedM) {
                mj.r = B64[1];
            }
#pragma unroll
            for (int r = 0; r < R; ++r) {
                f2 m = mj;
                if (MASK) {
                    float m0 = 1.f, m1 = 1.f;
                    if (!EQM) f2_split(mj, m0, m1);
                    const long long j0 = j_first + 2 * u;
                    m = f2_make(j0 == ti[r] ? 0.f : m0, j0 + 1 == ti[r] ? 0.f : m1);
                }
                t[r].template interact<MASK>(xj, yj, zj, m, e2);
            }
        }
    } else {
#pragma unroll
        for (int u = 0; u < CH / 2; ++u) {
            const float4 A = rec[2 * u];
            const float4 B = rec[2 * u + 1];
#pragma unroll
            for (int r = 0; r < R; ++r) {
                float m0 = B.z, m1 = B.w;
                if (MASK) {
                    const long long j0 = j_first + 2 * u;
                    if (j0 == ti[r]) m0 = 0.f;
                    if (j0 + 1 == ti[r]) m1 = 0.f;
                }
                t[r].interact(A, B, m0, m1, eps2);
            }
        }
    }
}

}  // namespace nb
