// epilogue.cuh — FP64 per-body epilogue of the force kernel: unit conversion, integrator
// kick/drift in the reference's exact operation order, the next generation's group-relative
// FP32 pack, and the per-body diagnostics terms. Every FP64 product/sum uses the
// non-contracting intrinsics so nvcc cannot fuse them into DFMA: given the same accelerations
// the update is bit-identical to the reference's -O2 (non-FMA) arithmetic.
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

__device__ __forceinline__ float extent_of(const double (&x)[3], bool live) {
    if (!live) return 0.f;
    const double e = fmax(fabs(x[0]), fmax(fabs(x[1]), fabs(x[2])));
    return __double2float_ru(e);  // never under-estimate
}

// ---- group pack ---------------------------------------------------------------------------
// One group = kGroupBodies consecutive stream rows handled by the kGroupBodies threads of a CTA
// (thread t holds row g0 + t). Bounding box of the attracting rows -> FP64 origin (box centre)
// and half-extents; every row is stored as hi = float((x - c) * 2^-e), lo = float of the rest.
// `sbox` is shared scratch of 6 floats per warp. All threads of the CTA must call this.
template <int NT>
__device__ __forceinline__ void pack_group(const PeerSet& out, double* const (&tx_out)[3], long long row,
                                           const double (&x)[3], float m, bool attracts, double inv_len,
                                           double inv_mass, float (*sbox)[6]) {
    static_assert(NT == kGroupBodies, "one thread per row of a group");
    const int tid = threadIdx.x;
    constexpr float kInf = __builtin_huge_valf();
    float bmin[3], bmax[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        bmin[k] = attracts ? __double2float_rd(x[k]) : kInf;
        bmax[k] = attracts ? __double2float_ru(x[k]) : -kInf;
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            bmin[k] = fminf(bmin[k], __shfl_xor_sync(0xffffffffu, bmin[k], o));
            bmax[k] = fmaxf(bmax[k], __shfl_xor_sync(0xffffffffu, bmax[k], o));
        }
    }
    __syncthreads();  // sbox may still be read from an earlier group
    if ((tid & 31) == 0) {
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            sbox[tid >> 5][k] = bmin[k];
            sbox[tid >> 5][3 + k] = bmax[k];
        }
    }
    __syncthreads();
    double c[3];
    float h[3];
    bool any = true;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        float lo = sbox[0][k], hi = sbox[0][3 + k];
#pragma unroll
        for (int w = 1; w < NT / 32; ++w) {
            lo = fminf(lo, sbox[w][k]);
            hi = fmaxf(hi, sbox[w][3 + k]);
        }
        any = any && (lo <= hi);
        // centre in FP64 (exact for two floats), clamped so an overflowing float bound cannot poison it
        c[k] = (lo <= hi) ? __dmul_rn(__dadd_rn((double)fmaxf(lo, -3.0e38f), (double)fminf(hi, 3.0e38f)), 0.5) : 0.0;
        const double half = (lo <= hi) ? fmax(__dadd_rn((double)hi, -c[k]), __dadd_rn(c[k], -(double)lo)) : 0.0;
        h[k] = __double2float_ru(__dmul_rn(half, inv_len));
    }
    double cs[3];  // the origin in packed units (exact: power-of-two scale) — what the readers subtract
#pragma unroll
    for (int k = 0; k < 3; ++k) cs[k] = __dmul_rn(c[k], inv_len);
    // rows dropped by the gather filter (and padding rows) sit at hi_x = 2^60 packed units with m = 0:
    // r^6 overflows to +inf, rsqrt(+inf) = +0, so they contribute exactly 0 even to the kernels that
    // do not multiply by the mass (uniform-mass variants)
    float fh[3], fl[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        const double d = __dadd_rn(__dmul_rn(x[k], inv_len), -cs[k]);
        const float h1 = (float)d;
        fh[k] = attracts ? h1 : (k == 0 ? kFarPacked : 0.f);
        fl[k] = attracts ? (float)__dadd_rn(d, -(double)h1) : 0.f;
    }
    const float fm = attracts ? (float)__dmul_rn((double)m, inv_mass) : 0.f;
    GroupHdr g;
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        g.c[k] = cs[k];
        g.h[k] = any ? h[k] : 0.f;
    }
    g.reach = any ? kNearFactor * fmaxf(h[0], fmaxf(h[1], h[2])) : 0.f;
    g.pad[0] = g.pad[1] = 0.f;
#pragma unroll
    for (int k = 0; k < 3; ++k) tx_out[k][row] = x[k];  // the row's FP64 position in stream order (this device only)
    const long long off = (row >> 1) * 8 + (row & 1);
    for (int q = 0; q < out.n; ++q) {  // this device, then every peer (stores over NVLink)
        float* rec = out.pk[q] + off;
        float* rlo = out.lo[q] + off;
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            rec[2 * k] = fh[k];
            rlo[2 * k] = fl[k];
        }
        rec[6] = fm;
        rlo[6] = 0.f;
        if (tid == 0) out.hdr[q][row / kGroupBodies] = g;
    }
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

// What the per-body epilogue hands to the group pack.
struct NextRow {
    double x[3];
    float m;
    bool attracts;
};

// Per-body epilogue for body row i. S[0..2]: packed acceleration sums, S[3]: packed potential sum.
// Modes that pack (kEuler, kVerletKickDrift) fill `nx` with the row's next position.
template <bool POT>
__device__ __forceinline__ void body_epilogue(const StepParams& p, long long i, const double (&S)[4],
                                              const Scale& cur, DiagTerms& dg, NextRow& nx) {
    const int mode = p.mode;
    const bool live = p.live[i] != 0;
    const bool pinned = p.pinned[i] != 0;
    const float m = p.mass[i];
    const double phi = POT ? __dmul_rn(S[3], cur.pot_unit) : 0.0;
    double x[3], v[3];
    if (mode == kDiagOnly) {  // stand-alone compute_diagnostics: every body, nothing written
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            x[k] = p.x[k][i];
            v[k] = p.v[k][i];
        }
        diag_accumulate(dg, x, v, m, phi, p.g, true);
        return;
    }
    double a[3], a_prev[3];
#pragma unroll
    for (int k = 0; k < 3; ++k) {
        // pinned bodies attract but get a = 0 (Physics.hpp:248-255); bodies dropped by the
        // gather filter keep their stale acceleration (never scattered to, :260)
        a_prev[k] = p.a[k][i];
        if (live) {
            a[k] = pinned ? 0.0 : __dmul_rn(S[k], cur.acc_unit);
            p.a[k][i] = a[k];
        } else {
            a[k] = a_prev[k];
        }
    }
    if (mode == kForceOnly && !POT) return;

#pragma unroll
    for (int k = 0; k < 3; ++k) {
        x[k] = p.x[k][i];
        v[k] = p.v[k][i];
    }
    if (mode == kForceOnly) {
        if (POT) diag_accumulate(dg, x, v, m, phi, p.g, live);
        return;
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
        // velocity Verlet kick, Physics.hpp:309-323. PrevAcceleration is what Acceleration held when this
        // substep's drift ran (:304): the value this pass is about to replace.
        if (!pinned) {
#pragma unroll
            for (int k = 0; k < 3; ++k) {
                p.a0[k][i] = a_prev[k];
                const double am = __dmul_rn(__dadd_rn(a_prev[k], a[k]), 0.5);
                v[k] = __dadd_rn(v[k], __dmul_rn(am, p.dt));
            }
            cap_speed(v, p.max_speed);
#pragma unroll
            for (int k = 0; k < 3; ++k) p.v[k][i] = v[k];
            // The reference's NEXT gravity system re-evaluates the gather filter with the kicked velocity
            // (Physics.hpp:45-48,198-199); the fused path skips that pass as long as nothing changed. A kick
            // that makes a velocity non-finite changes the filter: tell the host (it re-evaluates like the reference).
            if (p.filter_flip && body_live(x, v, m) != live) *p.filter_flip = 1;
        }
        if (POT) diag_accumulate(dg, x, v, m, phi, p.g, live);
        if (mode == kVerletKick) return;
        // drift of the NEXT substep, Physics.hpp:297-303, into the shadow positions: x itself stays the
        // position this frame ends with until the host commits the drift (same dt, nothing edited)
#pragma unroll
        for (int k = 0; k < 3; ++k) {
            if (!pinned) x[k] = __dadd_rn(x[k], __dadd_rn(__dmul_rn(v[k], p.dt), __dmul_rn(a[k], p.half_dt2)));
            p.xs[k][i] = x[k];
        }
    }
    // the gather filter the next pass applies (new positions, new velocities), and the row to pack
    const bool live_next = body_live(x, v, m);
    p.live_next[i] = live_next ? 1 : 0;
#pragma unroll
    for (int k = 0; k < 3; ++k) nx.x[k] = x[k];
    nx.m = m;
    nx.attracts = live_next;
}

}  // namespace nb
