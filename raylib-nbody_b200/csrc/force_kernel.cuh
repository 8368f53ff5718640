// force_kernel.cuh — grid/CTA structure of the force kernel (see force.cuh for the math).
//
// Grid: one CTA per (target tile, source segment). A target tile is NT*R consecutive
// STREAM rows; warp w owns the 32*R consecutive rows row0 + w*32*R + r*32 + lane, so in Hilbert order
// a warp's targets are one compact clump and few source groups are "near" it (`order` maps a stream
// row to its body row in the FP64 state, which stays in upload order); the source range of the launch is split into
// `nseg` tile-aligned segments so the grid can be sized to a whole number of waves of 148 SMs for
// any N. With nseg > 1 each CTA leaves FP64 partial sums in `part`; the last CTA to arrive for
// a target tile adds them in fixed segment order (bit-reproducible regardless of arrival order)
// and runs the epilogue.
//
// The j-stream reaches shared memory through a kStages-deep ring of 1-D TMA bulk copies
// (cp.async.bulk + mbarrier complete_tx): one elected thread issues an 8 KB copy of the tile's
// pair records plus a 192 B copy of its four group headers, nobody spends issue slots on loads,
// and every warp reads sources with broadcast LDS.128. A stage is handed back with a CTA-wide
// barrier per tile (NB_RING_SYNC 1, the shipped setting; per-warp `empty` mbarriers measured 4 % slower).
//
// Per source group (128 rows) every thread re-bases its R targets on the group's FP64 origin
// (positions of the CTA's targets sit in shared memory as FP64) and decides, warp-uniformly, between
// the one-float path and the two-float near path (force.cuh).
#pragma once
#include "epilogue.cuh"
#include "force.cuh"

#ifndef NB_FLUSH_PER_GROUP
#define NB_FLUSH_PER_GROUP 1  // 1: fold the FP32 sums into FP64 once per 128-source group instead of per 64-source chunk
#endif

namespace nb {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

// NU: unroll multiplier of the near path. 1 for the large-N variants (the near path is rare there and its code
// must not disturb the hot loop's schedule); 4 for the "mid" variants the planner picks when the world has few
// source tiles and a large share of the (warp, group) visits is near.
template <int NT, int R, bool POT, int MINB, int CH = kFlushChunk, bool EQM = false, bool D2 = false,
          bool PONLY = false, int NU = 1>
__global__ void __launch_bounds__(NT, MINB)
force_kernel(const __grid_constant__ StepParams p) {
    static_assert(NT == kGroupBodies, "a target slice r of a CTA is exactly one group of the next generation");
    static_assert(kGroupBodies % CH == 0, "flush chunks must not straddle groups");
    constexpr int IT = NT * R;
    constexpr int kRecPerTile = kTileBodies;  // float4 records per tile (2 per body pair)
    constexpr uint32_t kTileBytes = kTileBodies * 16;
    constexpr int kWarps = NT / 32;
    constexpr int kChunksPerGroup = kGroupBodies / CH;
    typedef Target<POT, EQM, D2, PONLY> Tg;

    __shared__ alignas(128) float4 ring[kStages][kRecPerTile];
    __shared__ alignas(16) GroupHdr ring_hdr[kStages][kGroupsPerTile];
    __shared__ alignas(8) uint64_t full_bar[kStages];
    __shared__ alignas(8) uint64_t empty_bar[kStages];
    __shared__ double tpos[3][IT];  // FP64 positions of this CTA's targets in packed units (x * 2^-e, exact);
                                    // re-used by the epilogue for the rows' next positions
    __shared__ float tmass[IT];
    __shared__ unsigned char tatt[IT];
    __shared__ double red[kWarps][kDiagValues];
    __shared__ float red_ext[kWarps];
    __shared__ float sbox[kWarps][6];
    __shared__ int is_last;

    const int tid = threadIdx.x;
    const long long unit = blockIdx.x;
    const long long it = unit / p.nseg;
    const int js = (int)(unit - it * p.nseg);
    const long long row0 = p.i_lo + it * IT;  // first stream row of the target tile
    const int li0 = (tid >> 5) * (32 * R) + (tid & 31);  // tile-local row of target r: li0 + 32 * r
    const long long rbase = row0 + li0;

    // scales of the generation being read and of the one the epilogue packs (common.cuh)
    Scale cur = scale_for(p.extent[p.slot_cur], p.mass_exp, p.g, p.softening);
    if (PONLY && !(p.softening > 0.f)) cur.eps2 = 0.f;  // compute_diagnostics: eps = 0 stays 0 (Physics.hpp:152)
    cur.acc_unit *= p.eqm_mass;
    cur.pot_unit *= p.eqm_mass;

    // ---- targets ---------------------------------------------------------------------
    // Rows past the rank's last target get a position no group can be "near" (their sums are never stored),
    // so the per-group loop below carries no validity test.
    Tg tg[R];
    double S[R][4];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const long long row = rbase + 32 * r;
        const bool ok = row < p.i_hi;
#pragma unroll
        for (int k = 0; k < 3; ++k) tpos[k][li0 + 32 * r] = ok ? __dmul_rn(p.tx_cur[k][row], cur.inv_len) : 1.0e300;
        tg[r].reset();
        S[r][0] = S[r][1] = S[r][2] = S[r][3] = 0.0;
    }

    // ---- source segment of this CTA (logical tiles) -----------------------------------
    const long long span = p.j_tile_hi - p.j_tile_lo;
    const long long t_begin = p.j_tile_lo + span * js / p.nseg;
    const long long t_end = p.j_tile_lo + span * (js + 1) / p.nseg;
    auto phys = [&](long long t) { return t >= p.hole_lo ? t + p.hole_len : t; };
    auto issue = [&](int s, long long t) {
        const long long pt = phys(t);
        mbar_arrive_expect_tx(&full_bar[s], kTileBytes + kHdrBytesPerTile);
        tma_load_1d(&ring[s][0], p.pk_cur + pt * (kTileBodies * 4), kTileBytes, &full_bar[s]);
        tma_load_1d(&ring_hdr[s][0], p.hdr_cur + pt * kGroupsPerTile, kHdrBytesPerTile, &full_bar[s]);
    };

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();  // barriers initialised, tpos filled
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s)
            if (t_begin + s < t_end) issue(s, t_begin + s);
    }

    for (long long t = t_begin; t < t_end; ++t) {
        const int k = (int)(t - t_begin);
        const int s = k % kStages;
        mbar_wait(&full_bar[s], (uint32_t)((k / kStages) & 1));
        const long long j_tile = phys(t) * kTileBodies;
        const bool diag_tile = POT && (j_tile < row0 + IT) && (j_tile + kTileBodies > row0);
#pragma unroll 1
        for (int g = 0; g < kGroupsPerTile; ++g) {
            // ---- re-base the targets on this group's origin; near test ---------------------
            const GroupHdr* gh = &ring_hdr[s][g];
            const double c0 = gh->c[0], c1 = gh->c[1], c2 = gh->c[2];
            const float4 hb = *reinterpret_cast<const float4*>(gh->h);  // {hx, hy, hz, reach}
            bool near = false;
#ifdef NB_EXP_GROUP_ONCE  // developer A/B only: what the per-group re-basing costs (wrong numerics)
            if (g == 0)
#endif
#pragma unroll
            for (int r = 0; r < R; ++r) {
                const float f0 = (float)__dadd_rn(tpos[0][li0 + 32 * r], -c0);
                const float f1 = (float)__dadd_rn(tpos[1][li0 + 32 * r], -c1);
                const float f2v = D2 ? 0.f : (float)__dadd_rn(tpos[2][li0 + 32 * r], -c2);
                tg[r].set_offset(f0, f1, f2v);
                const bool inside = (fabsf(f0) < hb.x + hb.w) && (fabsf(f1) < hb.y + hb.w) &&
                                    (D2 || fabsf(f2v) < hb.z + hb.w);
                near = near || inside;
            }
            const float4* grec = &ring[s][g * kGroupBodies];
            const long long j_group = j_tile + (long long)g * kGroupBodies;
#ifdef NB_EXP_NO_NEAR  // developer A/B only: what the near path costs (wrong numerics for near pairs)
            near = false;
#endif
            const bool near_any = __any_sync(0xffffffffu, near);
            if (!near_any) {
#pragma unroll 1
                for (int c = 0; c < kChunksPerGroup; ++c) {
                    const float4* rec = grec + c * CH;
                    const long long j_first = j_group + c * CH;
                    if (diag_tile)
                        chunk_interact<R, POT, true, CH, EQM, D2, PONLY, 32>(tg, rec, cur.eps2, j_first, rbase);
                    else
                        chunk_interact<R, POT, false, CH, EQM, D2, PONLY, 32>(tg, rec, cur.eps2, j_first, rbase);
#if !NB_FLUSH_PER_GROUP
#pragma unroll
                    for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
#endif
                }
#if NB_FLUSH_PER_GROUP
#pragma unroll
                for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
#endif
            } else {
                // ---- near group: two-float offsets on both sides, RN targets at a time ------------
                // (developer counter of nb_near_stats: kept out of the common path, where a lane-divergent
                // statement would cost the hot loop its uniform-register addressing)
                if (p.stats && (tid & 31) == 0) atomicAdd(&p.stats[1], 1ull);
#ifndef NB_NEAR_RN
#define NB_NEAR_RN 4  // targets per pass of the near path (4 + unroll 2 measured best: profiles/r02_time_variants.txt)
#endif
                constexpr int RN = R >= NB_NEAR_RN ? NB_NEAR_RN : (R >= 2 ? 2 : 1);
                const float4* glo = reinterpret_cast<const float4*>(p.lo_cur) + j_group;
#pragma unroll 1
                for (int c = 0; c < kChunksPerGroup; ++c) {
                    const float4* rec = grec + c * CH;
                    const float4* lrec = glo + c * CH;
                    const long long j_first = j_group + c * CH;
#pragma unroll
                    for (int r0 = 0; r0 < R; r0 += RN) {
                        if (r0 + RN > R) break;
                        f2 nl[RN][3];
#pragma unroll
                        for (int q = 0; q < RN; ++q) {
#pragma unroll
                            for (int k = 0; k < 3; ++k) {
                                const double ck = k == 0 ? c0 : (k == 1 ? c1 : c2);
                                const double d = __dadd_rn(tpos[k][li0 + 32 * (r0 + q)], -ck);
                                const float fh = (float)d;
                                const float fl = -(float)__dadd_rn(d, -(double)fh);
                                nl[q][k] = f2_make(fl, fl);
                            }
                        }
                        if (diag_tile)
                            chunk_interact_near<RN, R, POT, true, CH, EQM, D2, PONLY, 32, NU>(tg, r0, rec, lrec, nl, cur.eps2,
                                                                                          j_first, rbase);
                        else
                            chunk_interact_near<RN, R, POT, false, CH, EQM, D2, PONLY, 32, NU>(tg, r0, rec, lrec, nl, cur.eps2,
                                                                                           j_first, rbase);
                    }
                    if (R % RN) {  // R = 3: the odd target on its own
                        constexpr int rl = R - 1;
                        f2 nl1[1][3];
#pragma unroll
                        for (int k = 0; k < 3; ++k) {
                            const double ck = k == 0 ? c0 : (k == 1 ? c1 : c2);
                            const double d = __dadd_rn(tpos[k][li0 + 32 * rl], -ck);
                            const float fh = (float)d;
                            const float fl = -(float)__dadd_rn(d, -(double)fh);
                            nl1[0][k] = f2_make(fl, fl);
                        }
                        if (diag_tile)
                            chunk_interact_near<1, R, POT, true, CH, EQM, D2, PONLY, 32, NU>(tg, rl, rec, lrec, nl1, cur.eps2,
                                                                                         j_first, rbase);
                        else
                            chunk_interact_near<1, R, POT, false, CH, EQM, D2, PONLY, 32, NU>(tg, rl, rec, lrec, nl1, cur.eps2,
                                                                                          j_first, rbase);
                    }
#pragma unroll
                    for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
                }
            }
        }
#if NB_RING_SYNC
        __syncthreads();  // every warp is done with stage s
        if (tid == 0 && t + kStages < t_end) issue(s, t + kStages);
#else
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty_bar[s]);
        if (tid == 0 && k >= 1 && t - 1 + kStages < t_end) {
            const int sp = (k - 1) % kStages;
            mbar_wait(&empty_bar[sp], (uint32_t)(((k - 1) / kStages) & 1));
            issue(sp, t - 1 + kStages);
        }
#endif
    }

    // ---- combine source segments (fixed order) ------------------------------------------
    constexpr int NC = POT ? 4 : 3;
    constexpr int C0 = PONLY ? 3 : 0;  // potential-only kernels carry S[3] alone
    if (p.nseg > 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (rbase + 32 * r >= p.i_hi) continue;
            const long long col = rbase + 32 * r - p.i_lo;
#pragma unroll
            for (int c = C0; c < NC; ++c) __stcg(&p.part[((long long)(1 + js) * 4 + c) * p.i_cap + col], S[r][c]);
        }
        __threadfence();
        __syncthreads();
        if (tid == 0) {
            const unsigned int prev = atomicAdd(&p.tile_counter[it], 1u);
            is_last = (prev == (unsigned int)(p.nseg - 1));
            if (is_last) p.tile_counter[it] = 0;  // ready for the next launch
        }
        __syncthreads();
        if (!is_last) return;
        __threadfence();
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (rbase + 32 * r >= p.i_hi) continue;
            const long long col = rbase + 32 * r - p.i_lo;
#pragma unroll
            for (int c = C0; c < NC; ++c) {
                double acc = 0.0;
                for (int sgm = 0; sgm < p.nseg; ++sgm) acc += __ldcg(&p.part[((long long)(1 + sgm) * 4 + c) * p.i_cap + col]);
                S[r][c] = acc;
            }
        }
    }
    if (p.accumulate) {  // partial sums left by an earlier launch over other sources
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (rbase + 32 * r >= p.i_hi) continue;
            const long long col = rbase + 32 * r - p.i_lo;
#pragma unroll
            for (int c = C0; c < NC; ++c) S[r][c] += __ldcg(&p.part[(long long)c * p.i_cap + col]);
        }
    }
    if (p.mode == kPartialOnly) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (rbase + 32 * r >= p.i_hi) continue;
            const long long col = rbase + 32 * r - p.i_lo;
#pragma unroll
            for (int c = C0; c < NC; ++c) p.part[(long long)c * p.i_cap + col] = S[r][c];
        }
        return;
    }

    // ---- epilogue -----------------------------------------------------------------------
    const Scale nxt = scale_for(p.extent[p.slot_nxt], p.mass_exp, p.g, p.softening);
    const double inv_mass = pow2d(-p.mass_exp);
    const bool packs = (p.mode == kEuler || p.mode == kVerletKickDrift);
    DiagTerms dg;
#pragma unroll
    for (int k = 0; k < kDiagValues; ++k) dg.v[k] = 0.0;
    float ext = 0.f;
    if (packs) __syncthreads();  // every warp is done reading tpos
#pragma unroll
    for (int r = 0; r < R; ++r) {
        NextRow nx;
        nx.x[0] = nx.x[1] = nx.x[2] = 0.0;
        nx.m = 0.f;
        nx.attracts = false;
        if (rbase + 32 * r < p.i_hi) body_epilogue<POT>(p, (long long)p.order[rbase + 32 * r], S[r], cur, dg, nx);
        if (packs) {  // hand the row's next position to the pack below (which groups rows by 128, not by warp)
#pragma unroll
            for (int k = 0; k < 3; ++k) tpos[k][li0 + 32 * r] = nx.x[k];
            tmass[li0 + 32 * r] = nx.m;
            tatt[li0 + 32 * r] = nx.attracts ? 1 : 0;
        }
    }
    if (packs) {
        __syncthreads();
        // slice q of the tile (rows row0 + q*NT + tid) is exactly one group of generation gen+1
#pragma unroll
        for (int q = 0; q < R; ++q) {
            if (row0 + (long long)q * NT >= p.row_cap) break;  // CTA-uniform
            const double nx3[3] = {tpos[0][q * NT + tid], tpos[1][q * NT + tid], tpos[2][q * NT + tid]};
            const bool att = tatt[q * NT + tid] != 0;
            ext = fmaxf(ext, extent_of(nx3, att));
            pack_group<NT>(p.out, p.tx_next, row0 + (long long)q * NT + tid, nx3, tmass[q * NT + tid], att,
                           nxt.inv_len, inv_mass, sbox);
        }
    }
    if (packs) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) ext = fmaxf(ext, __shfl_xor_sync(0xffffffffu, ext, o));
        if ((tid & 31) == 0) red_ext[tid >> 5] = ext;
    }
    if (POT) {
#pragma unroll
        for (int k = 0; k < 10; ++k) {
            const double w = warp_sum(dg.v[k]);
            if ((tid & 31) == 0) red[tid >> 5][k] = w;
        }
    }
    __syncthreads();
    if (tid == 0 && packs) {
        float e = 0.f;
        for (int w = 0; w < kWarps; ++w) e = fmaxf(e, red_ext[w]);
        for (int q = 0; q < p.out.n; ++q) atomicMax(&p.out.extent[q][(p.gen + 1) % kExtentSlots], __float_as_uint(e));
        // recycled by the one CTA that always reaches this epilogue for target tile 0 (with nseg > 1 that is
        // the last arrival of the tile's segments, whichever segment it carries); own copy only
        if (it == 0) p.out.extent[0][(p.gen + 2) % kExtentSlots] = 0u;
    }
    if (POT && tid < 10) {
        double acc = 0.0;
        for (int w = 0; w < kWarps; ++w) acc += red[w][tid];
        p.diag_part[it * kDiagValues + tid] = acc;
    }
}

}  // namespace nb
