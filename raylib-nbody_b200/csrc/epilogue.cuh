// epilogue.cuh — FP64 per-body epilogue of the force kernel: unit conversion, integrator
// kick/drift in the reference's exact operation order, the next generation's FP32 pack,
// and the per-body diagnostics terms. Every FP64 product/sum uses the non-contracting
// intrinsics so nvcc cannot fuse them into DFMA: given the same accelerations the update
// is bit-identical to the reference's -O2 (non-FMA) arithmetic.
#pragma once
#include "common.cuh"

namespace nb {

__device__ __forceinline__ bool finite_d(double v) { return fabs(v) <= 1.7976931348623157e308; }

// Gather filter of Physics::compute_gravity, src/systems/Physics.hpp:198-199.
__device__ __forceinline__ bool body_live(const double (&x)[3], const double (&v)[3], float m) {
    return finite_d(x[0]) && finite_d(x[1]) && finite_d(x[2]) && finite_d(v[0]) && finite_d(v[1]) &&
           finite_d(v[2]) && m > 0.0f && m <= 3.4028234663852886e38f;
}

// Speed cap, Physics.hpp:280-287 and :315-322.
__device__ __forceinline__ void cap_speed(double (&v)[3], float max_speed) {
    if (max_speed > 0.0f) {
        double s2 = __dadd_rn(__dmul_rn(v[0], v[0]), __dmul_rn(v[1], v[1]));
        s2 = __dadd_rn(s2, __dmul_rn(v[2], v[2]));  // z == 0 in 2-D: adds +0.0 exactly
        const double vlen = __dsqrt_rn(s2);
        if (vlen > (double)max_speed) {
            const double s = __ddiv_rn((double)max_speed, vlen);
            v[0] = __dmul_rn(v[0], s);
            v[1] = __dmul_rn(v[1], s);
            v[2] = __dmul_rn(v[2], s);
        }
    }
}

// Write body i of a pair-SoA j-stream.
__device__ __forceinline__ void pack_body(float* __restrict__ pk, long long i, const double (&x)[3], float m,
                                          bool live, double inv_len, double inv_mass) {
    float* rec = pk + (i >> 1) * 8 + (i & 1);
    // bodies dropped by the gather filter (and padding rows) sit at x = 2^60 packed units with m = 0:
    // r^6 overflows to +inf, rsqrt(+inf) = +0, so they contribute exactly 0 even to the kernels
    // that do not multiply by the mass (uniform-mass variant)
    rec[0] = live ? (float)__dmul_rn(x[0], inv_len) : kFarPacked;
    rec[2] = live ? (float)__dmul_rn(x[1], inv_len) : 0.f;
    rec[4] = live ? (float)__dmul_rn(x[2], inv_len) : 0.f;
    rec[6] = live ? (float)__dmul_rn((double)m, inv_mass) : 0.f;
}

__device__ __forceinline__ float extent_of(const double (&x)[3], bool live) {
    if (!live) return 0.f;
    const double e = fmax(fabs(x[0]), fmax(fabs(x[1]), fabs(x[2])));
    return __double2float_ru(e);  // never under-estimate
}

struct DiagTerms {
    double v[kDiagValues];
};

// Diagnostics terms of one body, Physics.hpp:131-139 and the body's half of the pair sum :148-154.
__device__ __forceinline__ void diag_accumulate(DiagTerms& d, const double (&x)[3], const double (&v)[3], float mf,
                                                double phi_phys, double g, bool has_phi) {
    const double m = (double)mf;
    double v2 = __dadd_rn(__dmul_rn(v[0], v[0]), __dmul_rn(v[1], v[1]));
    v2 = __dadd_rn(v2, __dmul_rn(v[2], v[2]));
    d.v[0] += __dmul_rn(__dmul_rn(0.5, m), v2);
    if (has_phi) d.v[1] += -0.5 * g * m * phi_phys;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        d.v[2 + k] += __dmul_rn(m, v[k]);
        d.v[5 + k] += __dmul_rn(m, x[k]);
    }
    d.v[8] += m;
    if (!(finite_d(x[0]) && finite_d(x[1]) && finite_d(x[2]) && finite_d(v[0]) && finite_d(v[1]) && finite_d(v[2]) &&
          finite_d(m)))
        d.v[9] += 1.0;
}

// Per-body epilogue. S[0..2]: packed acceleration sums, S[3]: packed potential sum.
// Returns the extent contribution of the body's next position (0 when nothing was packed).
template <bool POT>
__device__ __forceinline__ float body_epilogue(const StepParams& p, long long i, const double (&S)[4],
                                               const Scale& cur, const Scale& nxt, double inv_mass, DiagTerms& dg) {
    const int mode = p.mode;
    const bool live = p.live[i] != 0;
    const bool pinned = p.pinned[i] != 0;
    const float m = p.mass[i];
    double a[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        // pinned bodies attract but get a = 0 (Physics.hpp:248-255); bodies dropped by the
        // gather filter keep their stale acceleration (never scattered to, :260)
        if (live) {
            a[k] = pinned ? 0.0 : __dmul_rn(S[k], cur.acc_unit);
            p.a[k][i] = a[k];
        } else {
            a[k] = p.a[k][i];
        }
    }
    if (mode == kForceOnly && !POT) return 0.f;

    double x[3], v[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        x[k] = p.x[k][i];
        v[k] = p.v[k][i];
    }
    const double phi = POT ? __dmul_rn(S[3], cur.pot_unit) : 0.0;
    if (mode == kForceOnly) {
        if (POT) diag_accumulate(dg, x, v, m, phi, p.g, live);
        return 0.f;
    }
    if (mode == kEuler) {
        // diagnostics describe the state the force pass saw (frame start), Physics.hpp:34-42
        if (POT) diag_accumulate(dg, x, v, m, phi, p.g, live);
        if (!pinned) {
#pragma unroll
            for (int k = 0; k < 3; ++k) v[k] = __dadd_rn(v[k], __dmul_rn(a[k], p.dt));  // :278-279
            cap_speed(v, p.max_speed);
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                x[k] = __dadd_rn(x[k], __dmul_rn(v[k], p.dt));  // :288-289
                p.v[k][i] = v[k];
                p.x[k][i] = x[k];
            }
        }
    } else {
        // velocity Verlet kick, Physics.hpp:309-323
        if (!pinned) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                const double am = __dmul_rn(__dadd_rn(p.a0[k][i], a[k]), 0.5);
                v[k] = __dadd_rn(v[k], __dmul_rn(am, p.dt));
            }
            cap_speed(v, p.max_speed);
#pragma unroll
            for (int k = 0; k < 3; ++k) p.v[k][i] = v[k];
        }
        if (POT) diag_accumulate(dg, x, v, m, phi, p.g, live);
        if (mode == kVerletKick) return 0.f;
        // drift of the next substep, Physics.hpp:297-304
        if (!pinned) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                x[k] = __dadd_rn(x[k], __dadd_rn(__dmul_rn(v[k], p.dt), __dmul_rn(a[k], p.half_dt2)));
                p.x[k][i] = x[k];
                p.a0[k][i] = a[k];
            }
        }
    }
    // pack generation gen+1 and re-evaluate the gather filter on the new state
    const bool live_next = body_live(x, v, m);
    p.live[i] = live_next ? 1 : 0;
    pack_body(p.pk_next, i, x, m, live_next, nxt.inv_len, inv_mass);
    return extent_of(x, live_next);
}

}  // namespace nb
