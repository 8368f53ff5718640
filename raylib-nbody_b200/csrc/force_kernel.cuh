// force_kernel.cuh — grid/CTA structure of the force kernel (see force.cuh for the math).
//
// Grid: one CTA per (target tile, source segment). A target tile is NT*R consecutive
// bodies (thread t owns bodies i0 + r*NT + t, so all FP64 state traffic is coalesced);
// the source range of the launch is split into `nseg` tile-aligned segments so the grid
// can be sized to a whole number of waves of 148 SMs for any N. With nseg > 1 each CTA
// leaves FP64 partial sums in `part`; the last CTA to arrive for a target tile adds them
// in fixed segment order (bit-reproducible regardless of arrival order) and runs the epilogue.
//
// The j-stream reaches shared memory through a kStages-deep ring of 1-D TMA bulk copies
// (cp.async.bulk + mbarrier complete_tx): one elected thread issues 8 KB copies, nobody
// spends issue slots on loads, and every warp reads sources with broadcast LDS.128. Stages are
// handed back through `empty` mbarriers (one arrival per warp), not a CTA-wide barrier.
#pragma once
#include "epilogue.cuh"
#include "force.cuh"

namespace nb {

__device__ __forceinline__ double warp_sum(double v) {
#pragma unroll
    for (int o = 16; o > 0; o >>= 1) v += __shfl_xor_sync(0xffffffffu, v, o);
    return v;
}

template <int NT, int R, bool PACKED, bool POT, int MINB, int CH = kFlushChunk, bool EQM = false, bool D2 = false>
__global__ void __launch_bounds__(NT, MINB)
force_kernel(const __grid_constant__ StepParams p) {
    constexpr int IT = NT * R;
    constexpr int kRecPerTile = kTileBodies;  // float4 records per tile (2 per body pair)
    constexpr uint32_t kTileBytes = kTileBodies * 16;
    constexpr int kWarps = NT / 32;

    __shared__ alignas(128) float4 ring[kStages][kRecPerTile];
    __shared__ alignas(8) uint64_t full_bar[kStages];
    __shared__ alignas(8) uint64_t empty_bar[kStages];
    __shared__ double red[kWarps][kDiagValues];
    __shared__ float red_ext[kWarps];
    __shared__ int is_last;

    const int tid = threadIdx.x;
    const long long unit = blockIdx.x;
    const long long it = unit / p.nseg;
    const int js = (int)(unit - it * p.nseg);
    const long long i0 = p.i_lo + it * IT;

    // scales of the generation being read and of the one the epilogue packs (common.cuh)
    Scale cur = scale_for(p.extent[(p.gen + kExtentSlots - 1) % kExtentSlots], p.mass_exp, p.g, p.softening);
    cur.acc_unit *= p.eqm_mass;
    cur.pot_unit *= p.eqm_mass;

    // ---- targets ---------------------------------------------------------------------
    Target<PACKED, POT, EQM, D2> tg[R];
    long long ti[R];
    double S[R][4];
#pragma unroll
    for (int r = 0; r < R; ++r) {
        const long long i = i0 + (long long)r * NT + tid;
        ti[r] = i < p.i_hi ? i : -1;
        const long long ic = i < p.i_hi ? i : p.i_hi - 1;
        const float* rec = p.pk_cur + (ic >> 1) * 8 + (ic & 1);
        tg[r].init(rec[0], rec[2], rec[4]);
        S[r][0] = S[r][1] = S[r][2] = S[r][3] = 0.0;
    }

    // ---- source segment of this CTA (logical tiles) -----------------------------------
    const long long span = p.j_tile_hi - p.j_tile_lo;
    const long long t_begin = p.j_tile_lo + span * js / p.nseg;
    const long long t_end = p.j_tile_lo + span * (js + 1) / p.nseg;
    auto phys = [&](long long t) { return t >= p.hole_lo ? t + p.hole_len : t; };

    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            mbar_init(&full_bar[s], 1);
            mbar_init(&empty_bar[s], kWarps);
        }
        fence_mbar_init();
    }
    __syncthreads();
    if (tid == 0) {
#pragma unroll
        for (int s = 0; s < kStages; ++s) {
            if (t_begin + s < t_end) {
                mbar_arrive_expect_tx(&full_bar[s], kTileBytes);
                tma_load_1d(&ring[s][0], p.pk_cur + phys(t_begin + s) * (kTileBodies * 4), kTileBytes, &full_bar[s]);
            }
        }
    }

    // No CTA-wide barrier in the tile loop: each warp releases a stage through its `empty`
    // mbarrier, and the producer thread refills the stage released one tile earlier, so it only
    // ever waits for a warp that has fallen a whole tile behind.
    for (long long t = t_begin; t < t_end; ++t) {
        const int k = (int)(t - t_begin);
        const int s = k % kStages;
        mbar_wait(&full_bar[s], (uint32_t)((k / kStages) & 1));
        const long long j_tile = phys(t) * kTileBodies;
        const bool diag_tile = POT && (j_tile < i0 + IT) && (j_tile + kTileBodies > i0);
#pragma unroll 1
        for (int c = 0; c < kTileBodies / CH; ++c) {
            const float4* rec = &ring[s][c * CH];
            const long long j_first = j_tile + c * CH;
            if (diag_tile)
                chunk_interact<R, PACKED, POT, true, CH, EQM, D2>(tg, rec, cur.eps2, j_first, ti);
            else
                chunk_interact<R, PACKED, POT, false, CH, EQM, D2>(tg, rec, cur.eps2, j_first, ti);
#pragma unroll
            for (int r = 0; r < R; ++r) tg[r].flush(S[r][0], S[r][1], S[r][2], S[r][3]);
        }
#if NB_RING_SYNC
        __syncthreads();  // every warp is done with stage s
        if (tid == 0 && t + kStages < t_end) {
            mbar_arrive_expect_tx(&full_bar[s], kTileBytes);
            tma_load_1d(&ring[s][0], p.pk_cur + phys(t + kStages) * (kTileBodies * 4), kTileBytes, &full_bar[s]);
        }
#else
        __syncwarp();
        if ((tid & 31) == 0) mbar_arrive(&empty_bar[s]);
        if (tid == 0 && k >= 1 && t - 1 + kStages < t_end) {
            const int sp = (k - 1) % kStages;
            mbar_wait(&empty_bar[sp], (uint32_t)(((k - 1) / kStages) & 1));
            mbar_arrive_expect_tx(&full_bar[sp], kTileBytes);
            tma_load_1d(&ring[sp][0], p.pk_cur + phys(t - 1 + kStages) * (kTileBodies * 4), kTileBytes, &full_bar[sp]);
        }
#endif
    }

    // ---- combine source segments (fixed order) ------------------------------------------
    constexpr int NC = POT ? 4 : 3;
    if (p.nseg > 1) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (ti[r] < 0) continue;
            const long long col = ti[r] - p.i_lo;
#pragma unroll
            for (int c = 0; c < NC; ++c) __stcg(&p.part[((long long)(1 + js) * 4 + c) * p.i_cap + col], S[r][c]);
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
            if (ti[r] < 0) continue;
            const long long col = ti[r] - p.i_lo;
#pragma unroll
            for (int c = 0; c < NC; ++c) {
                double acc = 0.0;
                for (int sgm = 0; sgm < p.nseg; ++sgm) acc += __ldcg(&p.part[((long long)(1 + sgm) * 4 + c) * p.i_cap + col]);
                S[r][c] = acc;
            }
        }
    }
    if (p.accumulate) {  // partial sums left by an earlier launch over other sources
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (ti[r] < 0) continue;
            const long long col = ti[r] - p.i_lo;
#pragma unroll
            for (int c = 0; c < NC; ++c) S[r][c] += __ldcg(&p.part[(long long)c * p.i_cap + col]);
        }
    }
    if (p.mode == kPartialOnly) {
#pragma unroll
        for (int r = 0; r < R; ++r) {
            if (ti[r] < 0) continue;
            const long long col = ti[r] - p.i_lo;
#pragma unroll
            for (int c = 0; c < NC; ++c) p.part[(long long)c * p.i_cap + col] = S[r][c];
        }
        return;
    }

    // ---- epilogue -----------------------------------------------------------------------
    const Scale nxt = scale_for(p.extent[p.gen % kExtentSlots], p.mass_exp, p.g, p.softening);
    const double inv_mass = pow2d(-p.mass_exp);
    DiagTerms dg;
#pragma unroll
    for (int k = 0; k < kDiagValues; ++k) dg.v[k] = 0.0;
    float ext = 0.f;
#pragma unroll
    for (int r = 0; r < R; ++r) {
        if (ti[r] < 0) continue;
        ext = fmaxf(ext, body_epilogue<POT>(p, ti[r], S[r], cur, nxt, inv_mass, dg));
    }
    const bool packs = (p.mode == kEuler || p.mode == kVerletKickDrift);
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
        atomicMax(&p.extent[(p.gen + 1) % kExtentSlots], __float_as_uint(e));
        // recycled by the one CTA that always reaches this epilogue for target tile 0 (with nseg > 1 that is
        // the last arrival of the tile's segments, whichever segment it carries)
        if (it == 0) p.extent[(p.gen + 2) % kExtentSlots] = 0u;
    }
    if (POT && tid < 10) {
        double acc = 0.0;
        for (int w = 0; w < kWarps; ++w) acc += red[w][tid];
        p.diag_part[it * kDiagValues + tid] = acc;
    }
}

}  // namespace nb
