// force.cuh — the hot kernel's arithmetic: j-tiled direct-sum softened gravity on pairs of sources.
//
// Replaces the reference's hot loop A, src/systems/Physics.hpp:231-258 (direct all-pairs
// branch of compute_gravity), and — in the potential-only variant — hot loop B, the pair sum of
// compute_diagnostics (Physics.hpp:148-161). It is a full N x N target-major sweep (each target
// sums all sources in ascending stream order), not the reference's triangular pair loop, so
// no atomics and a fixed, reproducible summation order.
//
// Per interaction (3-D): 3 FADD (dx), 3 FFMA (r^2 seeded with eps^2), 2 FMUL (r^6), 1 MUFU.RSQ
// (r^-3 directly: one rsqrt error instead of three), 1 FMUL (m r^-3), 3 FFMA  = 12 FP32-pipe
// ops + 1 SFU op = the conventional 20 FLOPs, issued as two-wide FFMA2/FADD2/FMUL2 on a pair of
// sources. Compute-bound on the FP32 pipe; HBM sees the j-stream once per CTA through L2 (TMA
// bulk copies into a shared-memory ring).
//
// Positions reach this code relative to the FP64 origin of the SOURCE group (common.cuh): the
// source's offset is in the stream, the target's offset (x_i - c_g) is formed in FP64 once per
// (target, group) and rounded to float, so dx = hi_j - hi_i carries a relative error of a few
// 2^-25 for every pair whose target is clear of the group's bounding box. Targets inside or next to
// the box ("near") add the float residuals of both sides: dx = (hi_j - hi_i) + (lo_j - lo_i),
// three FADD2 per axis instead of one — the reference's FP64 `dx = pos[j] - pos[i]`
// (Physics.hpp:237-238) to 2^-48 of the group's extent before it is rounded to float.
#pragma once
#include "common.cuh"
#include "ptx.cuh"

#ifndef NB_UNROLL_PAIRS
#define NB_UNROLL_PAIRS 32  // pair records per unrolled loop body (32 = a whole 64-source flush chunk)
#endif

namespace nb {

constexpr int kUnrollPairs = NB_UNROLL_PAIRS;
#ifndef NB_NEAR_UNROLL
#define NB_NEAR_UNROLL 2
#endif
constexpr int kNearUnroll = NB_NEAR_UNROLL;  // pair records per loop body of the near path at 3-4 targets per pass
                                             // (x2 at two targets, x4 at one: small worlds run R = 1 and are all "near")

// float -> double of a chunk sum. F2F.F64.F32 runs on the XU pipe next to the kernel's MUFU.RSQ stream
// (one conversion costs as much XU time as a MUFU); NB_INT_FLUSH builds the double from the float's bits
// with integer-pipe instructions instead (exact for normal floats and +-0; chunk sums are never denormal
// under .ftz; inf/nan keep their class).
__device__ __forceinline__ double widen(float f) {
#ifdef NB_INT_FLUSH
    const unsigned int u = __float_as_uint(f);
    const unsigned int mag = u & 0x7fffffffu;
    unsigned int hi = (mag >> 3) + 0x38000000u;       // rebias the exponent: 1023 - 127 = 896 = 0x380
    hi = mag == 0u ? 0u : hi;
    hi = mag >= 0x7f800000u ? (0x7ff00000u | (mag >> 3)) : hi;  // inf / nan
    return __hiloint2double((int)(hi | (u & 0x80000000u)), (int)(u << 29));
#else
    return (double)f;
#endif
}

// One target body's running sums: FP32 inside a flush chunk, FP64 across chunks.
// EQM: all sources carry the same mass -> the m*w multiply is dropped (the common factor is
// applied in FP64 in the epilogue); D2: z == 0 for every body (reference 2-D uploads);
// PONLY: only the potential sum (stand-alone compute_diagnostics).
template <bool POT, bool EQM = false, bool D2 = false, bool PONLY = false>
struct Target {
    f2 nxi, nyi, nzi;  // {-hi_i, -hi_i} relative to the current source group's origin
    f2 ax, ay, az, ph;

    __device__ __forceinline__ void reset() { ax = ay = az = ph = f2_make(0.f, 0.f); }
    __device__ __forceinline__ void set_offset(float x, float y, float z) {
        nxi = f2_make(-x, -x); nyi = f2_make(-y, -y); nzi = f2_make(-z, -z);
    }

    // mj: the sources' masses; for EQM kernels only read when MASKED, as a {0|1, 0|1} self-term mask.
    // bump (PONLY && MASKED): {1|0, 1|0} added to r^2 of the self term so eps = 0 gives 0 * finite, not 0 * inf.
    template <bool MASKED>
    __device__ __forceinline__ void finish(f2 dx, f2 dy, f2 dz, f2 mj, f2 eps2, f2 bump) {
        f2 r2 = f2_fma(dx, dx, eps2);
        r2 = f2_fma(dy, dy, r2);
        if (!D2) r2 = f2_fma(dz, dz, r2);
        float lo, hi;
        if (PONLY) {
            if (MASKED) r2 = f2_add(r2, bump);
            f2_split(r2, lo, hi);
            const f2 w = f2_make(rsqrt_fast(lo), rsqrt_fast(hi));  // 1 / r
            if (EQM && !MASKED) ph = f2_add(w, ph);
            else ph = f2_fma(mj, w, ph);
            return;
        }
        const f2 r6 = f2_mul(f2_mul(r2, r2), r2);
        f2_split(r6, lo, hi);
        const f2 w = f2_make(rsqrt_fast(lo), rsqrt_fast(hi));  // r^-3
        const f2 s = EQM ? w : f2_mul(mj, w);
        ax = f2_fma(s, dx, ax);
        ay = f2_fma(s, dy, ay);
        if (!D2) az = f2_fma(s, dz, az);
        if (POT) {
            if (EQM && !MASKED) ph = f2_fma(w, r2, ph);  // 1 / r (non-attracting rows: w = 0)
            else ph = f2_fma(mj, f2_mul(w, r2), ph);    // m / r
        }
    }

    template <bool MASKED>
    __device__ __forceinline__ void interact(f2 xj, f2 yj, f2 zj, f2 mj, f2 eps2, f2 bump) {
        const f2 dx = f2_add(xj, nxi), dy = f2_add(yj, nyi);
        f2 dz{0ull};
        if (!D2) dz = f2_add(zj, nzi);
        finish<MASKED>(dx, dy, dz, mj, eps2, bump);
    }

    // near path: (hi_j - hi_i) + (lo_j - lo_i); nxl.. = {-lo_i, -lo_i}
    template <bool MASKED>
    __device__ __forceinline__ void interact_near(f2 xj, f2 yj, f2 zj, f2 mj, f2 xl, f2 yl, f2 zl, f2 nxl, f2 nyl,
                                                  f2 nzl, f2 eps2, f2 bump) {
        const f2 dx = f2_add(f2_add(xj, nxi), f2_add(xl, nxl));
        const f2 dy = f2_add(f2_add(yj, nyi), f2_add(yl, nyl));
        f2 dz{0ull};
        if (!D2) dz = f2_add(f2_add(zj, nzi), f2_add(zl, nzl));
        finish<MASKED>(dx, dy, dz, mj, eps2, bump);
    }

    // The even/odd-source halves are added in FP32 first: one commutative, sign-symmetric
    // rounding, so mirror-symmetric configurations (the reference's three-body seed,
    // Physics.hpp:87-115) keep exactly mirrored sums whatever slot a source landed in.
    __device__ __forceinline__ void flush(double& Ax, double& Ay, double& Az, double& Ph) {
        float lo, hi;
        if (!PONLY) {
            f2_split(ax, lo, hi); Ax += widen(lo + hi);
            f2_split(ay, lo, hi); Ay += widen(lo + hi);
            if (!D2) { f2_split(az, lo, hi); Az += widen(lo + hi); }
            ax = ay = az = f2_make(0.f, 0.f);
        }
        if (POT) { f2_split(ph, lo, hi); Ph += widen(lo + hi); ph = f2_make(0.f, 0.f); }
    }
};

// Load one pair record {x0,x1,y0,y1 | z0,z1,m0,m1} from a float4 pair (shared or global memory).
template <bool NEED_Z, bool NEED_M>
__device__ __forceinline__ void load_pair(const float4* __restrict__ rec, f2& xj, f2& yj, f2& zj, f2& mj) {
    const ulonglong2 A = *reinterpret_cast<const ulonglong2*>(rec);
    xj.r = A.x;
    yj.r = A.y;
    zj.r = 0ull;
    mj.r = 0ull;
    const unsigned long long* B64 = reinterpret_cast<const unsigned long long*>(rec + 1);
    if (NEED_Z && NEED_M) {
        const ulonglong2 B = *reinterpret_cast<const ulonglong2*>(rec + 1);
        zj.r = B.x;
        mj.r = B.y;
    } else if (NEED_Z) {
        zj.r = B64[0];
    } else if (NEED_M) {
        mj.r = B64[1];
    }
}

// Self-term mask of a pair record for the target at stream row `row`: its own mass (and, PONLY, its r^2 bump).
template <bool EQM>
__device__ __forceinline__ void self_mask(f2 mj, long long j0, long long row, f2& m, f2& bump) {
    float m0 = 1.f, m1 = 1.f;
    if (!EQM) f2_split(mj, m0, m1);
    const bool s0 = j0 == row, s1 = j0 + 1 == row;
    m = f2_make(s0 ? 0.f : m0, s1 ? 0.f : m1);
    bump = f2_make(s0 ? 1.f : 0.f, s1 ? 1.f : 0.f);
}

// One flush chunk (CH sources = CH/2 pair records) against R targets, hi parts only.
// MASK: the chunk may contain one of this thread's own targets (stream rows row0 + r*RS); its mass is
// zeroed for that target so the potential carries no self term (the force self term is exactly 0 anyway).
template <int R, bool POT, bool MASK, int CH, bool EQM, bool D2, bool PONLY, int RS>
__device__ __forceinline__ void chunk_interact(Target<POT, EQM, D2, PONLY> (&t)[R], const float4* __restrict__ rec,
                                               float eps2, long long j_first, long long row0) {
    const f2 e2 = f2_make(eps2, eps2);
    const f2 zero = f2_make(0.f, 0.f);
    if constexpr (MASK) {
        // the one tile of a launch that holds the CTA's own targets: kept rolled (the per-pair self masks
        // of a fully unrolled body cost more registers than the whole hot loop)
#pragma unroll 2
        for (int u = 0; u < CH / 2; ++u) {
            f2 xj, yj, zj, mj;
            load_pair<!D2, !EQM>(rec + 2 * u, xj, yj, zj, mj);
#pragma unroll
            for (int r = 0; r < R; ++r) {
                f2 m, bump;
                self_mask<EQM>(mj, j_first + 2 * u, row0 + (long long)r * RS, m, bump);
                t[r].template interact<true>(xj, yj, zj, m, e2, bump);
            }
        }
    } else {
#pragma unroll kUnrollPairs
        for (int u = 0; u < CH / 2; ++u) {
            f2 xj, yj, zj, mj;
            load_pair<!D2, !EQM>(rec + 2 * u, xj, yj, zj, mj);
#pragma unroll
            for (int r = 0; r < R; ++r) t[r].template interact<false>(xj, yj, zj, mj, e2, zero);
        }
    }
}

// The same chunk for RN consecutive targets t[r0 .. r0+RN) through the two-float path. lo_rec: the chunk's
// residual records in global memory (L2/L1-resident; every lane reads the same address). nl[q][k]:
// {-lo_i, -lo_i} of target r0+q, axis k. Partially unrolled: this path runs for the handful of groups
// whose bounding box a warp's targets touch, and must not cost the hot loop registers.
template <int RN, int R, bool POT, bool MASK, int CH, bool EQM, bool D2, bool PONLY, int RS, int NU = 1>
__device__ __forceinline__ void chunk_interact_near(Target<POT, EQM, D2, PONLY> (&t)[R], int r0,
                                                    const float4* __restrict__ rec,
                                                    const float4* __restrict__ lo_rec, const f2 (&nl)[RN][3],
                                                    float eps2, long long j_first, long long row0) {
    const f2 e2 = f2_make(eps2, eps2);
    const f2 zero = f2_make(0.f, 0.f);
#pragma unroll (NU * (RN >= 3 ? kNearUnroll : (RN == 2 ? 2 * kNearUnroll : 4 * kNearUnroll)))
    for (int u = 0; u < CH / 2; ++u) {
        f2 xj, yj, zj, mj, xl, yl, zl, unused;
        load_pair<!D2, !EQM>(rec + 2 * u, xj, yj, zj, mj);
        load_pair<!D2, false>(lo_rec + 2 * u, xl, yl, zl, unused);
#pragma unroll
        for (int q = 0; q < RN; ++q) {
            if (MASK) {
                f2 m, bump;
                self_mask<EQM>(mj, j_first + 2 * u, row0 + (long long)(r0 + q) * RS, m, bump);
                t[r0 + q].template interact_near<true>(xj, yj, zj, m, xl, yl, zl, nl[q][0], nl[q][1], nl[q][2], e2, bump);
            } else {
                t[r0 + q].template interact_near<false>(xj, yj, zj, mj, xl, yl, zl, nl[q][0], nl[q][1], nl[q][2], e2, zero);
            }
        }
    }
}

}  // namespace nb
