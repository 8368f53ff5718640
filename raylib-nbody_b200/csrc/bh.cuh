// bh.cuh — the reference's Barnes-Hut branch (src/systems/Physics.hpp:212-230 over the quadtree of
// src/physics/SpatialPartition.hpp:36-220), bit-identical, on the device.
//
// With its default settings the reference leaves the all-pairs loop as soon as a world has more
// than bh_threshold = 100 bodies (Constants.hpp:90) and walks a float point-region quadtree instead.
// A drop-in for the reference's frame therefore needs this branch as well as the direct sum. What
// makes the result reproducible bit for bit on a GPU:
//
//   * the SHAPE of the tree depends only on the set of float positions (a cell that holds two or
//     more bodies has four children, one body per leaf), not on insertion order, so it can be grown
//     breadth-first: one level per iteration, every body stepping down one cell per level, inside
//     ONE persistent cooperative launch with grid-wide barriers between the levels;
//   * cell centres are produced by the same float operations along the same root-to-leaf path
//     (c -+ half/2), the quadrant test is the same pair of float compares (:169-176);
//   * masses and centres of mass are float sums over the four children in the order 0..3 (:201-212):
//     one thread per cell, deepest level first;
//   * the walk is depth-first with the children visited 3,2,1,0 (a LIFO stack filled 0..3, :100-103)
//     and its terms are rounded to float and added to a float accumulator in that order (:83-98).
//     Each body's walk is sequential on one thread; a skip link per cell (next cell in that order
//     when the subtree is not opened) replaces the stack. The accept test (2*half)/|d| < theta is
//     decided in FP32 when the quotient is clear of theta by more than 1e-5 (relative; the FP32
//     evaluation is good to 1e-6) and re-evaluated exactly as the reference does in FP64 otherwise,
//     so every decision equals the reference's; accepted terms use the reference's FP64 sequence.
//   * threads are assigned to bodies in Z-order of their first 16 tree levels (one radix sort per
//     pass), so the lanes of a warp walk nearly the same cells: loads coalesce into a few sectors of
//     the L2-resident tree and the warp's trip count is close to each lane's own.
//
// Two bodies with identical float coordinates never separate; the reference then subdivides until
// it runs out of memory. Here the build stops at kBhMaxLevels and reports the condition.
#pragma once
#include <cooperative_groups.h>

#include "state.cuh"

namespace nb {

namespace cg = cooperative_groups;

constexpr int kBhMaxLevels = 320;  // a float half-size underflows to 0 after < 300 halvings
constexpr int kBhBuildThreads = 256;
#ifndef NB_BH_WALK_NT
#define NB_BH_WALK_NT 128
#endif
#ifndef NB_BH_WALK_MINB
#define NB_BH_WALK_MINB 1
#endif
constexpr int kBhWalkThreads = NB_BH_WALK_NT;
constexpr int kBhKeyLevels = 16;   // tree levels in the Z-order key
constexpr int kBhHistCells = 4096; // levels with at most this many new cells count bodies per CTA first

// control words (device, mirrored to pinned host memory after the build)
enum BhCtl { kBhNodes = 0, kBhError, kBhLive, kBhDepth, kBhMinX, kBhMaxX, kBhMinY, kBhMaxY, kBhCtlWords };
enum BhErr { kBhOk = 0, kBhOverflow = 1, kBhTooDeep = 2 };

struct BhTree {
    float2* bpos;        // float(x), float(y) per body row (Physics.hpp:217)
    float2* centre;      // cell centre (build only)
    float4* cell;        // {com.x, com.y, mass, half}
    int2* link;          // x: first child (>= 0) | -1 empty leaf | -2 - body for a leaf with a body; y: skip link
    int* count;          // bodies inside the cell (build only)
    int* cur;            // per body: cell it currently sits in, -1 once placed / not in the tree
    unsigned int* key;   // per body: Z-order key of its path
    int* order_in;       // iota
    int* level_begin;    // [kBhMaxLevels + 2]
    unsigned int* ctl;   // BhCtl
    int capacity;        // cells allocated
    unsigned long long* trace;  // developer knob NB_BH_TRACE: globaltimer stamps of the build's phases, else null
};

__device__ __forceinline__ void bh_stamp(const BhTree& t, long long gid, int& slot) {
    if (gid == 0 && t.trace) {
        unsigned long long now;
        asm volatile("mov.u64 %0, %%globaltimer;" : "=l"(now));
        t.trace[slot] = now;
    }
    ++slot;
}

__device__ __forceinline__ unsigned int float_order_key(float f) {
    const unsigned int u = __float_as_uint(f);
    return (u & 0x80000000u) ? ~u : (u | 0x80000000u);
}
__device__ __forceinline__ float float_from_order_key(unsigned int k) {
    return __uint_as_float((k & 0x80000000u) ? (k & 0x7fffffffu) : ~k);
}

// SpatialPartition.hpp:169-176
__device__ __forceinline__ int bh_quadrant(float2 c, float2 p) { return (p.x > c.x ? 1 : 0) + (p.y > c.y ? 2 : 0); }

// Build + aggregate in one cooperative launch (grid = resident CTAs only).
__global__ void __launch_bounds__(kBhBuildThreads) bh_build_kernel(StateView s, BhTree t) {
    cg::grid_group grid = cg::this_grid();
    __shared__ int s_hist[kBhHistCells];
    const long long gsz = (long long)gridDim.x * blockDim.x;
    const long long gid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    const int n = (int)s.n;
    int stamp = 1;
    bh_stamp(t, gid, stamp);

    // ---- gather filter (Physics.hpp:197-205), float casts (:217), bounding box (SpatialPartition.hpp:40-49)
    if (gid == 0) {
        t.ctl[kBhNodes] = 1; t.ctl[kBhError] = kBhOk; t.ctl[kBhLive] = 0; t.ctl[kBhDepth] = 0;
        t.ctl[kBhMinX] = 0xffffffffu; t.ctl[kBhMaxX] = 0u; t.ctl[kBhMinY] = 0xffffffffu; t.ctl[kBhMaxY] = 0u;
        t.link[0] = make_int2(-1, -1);
        t.count[0] = 0;
    }
    grid.sync();
    {
        unsigned int lo_x = 0xffffffffu, hi_x = 0u, lo_y = 0xffffffffu, hi_y = 0u, cnt = 0;
        for (long long i = gid; i < n; i += gsz) {
            const double x[3] = {s.x[0][i], s.x[1][i], 0.0};
            const double v[3] = {s.v[0][i], s.v[1][i], 0.0};
            const bool live = body_live(x, v, s.mass[i]);
            s.live[i] = live ? 1 : 0;
            t.order_in[i] = (int)i;
            if (!live) {
                t.cur[i] = -1;
                t.key[i] = 0xffffffffu;  // sorted behind every body of the tree
                continue;
            }
            const float2 p = make_float2(__double2float_rn(x[0]), __double2float_rn(x[1]));
            t.bpos[i] = p;
            t.cur[i] = 0;
            t.key[i] = 0u;
            const unsigned int kx = float_order_key(p.x), ky = float_order_key(p.y);
            lo_x = min(lo_x, kx); hi_x = max(hi_x, kx);
            lo_y = min(lo_y, ky); hi_y = max(hi_y, ky);
            ++cnt;
        }
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            lo_x = min(lo_x, __shfl_xor_sync(0xffffffffu, lo_x, o));
            hi_x = max(hi_x, __shfl_xor_sync(0xffffffffu, hi_x, o));
            lo_y = min(lo_y, __shfl_xor_sync(0xffffffffu, lo_y, o));
            hi_y = max(hi_y, __shfl_xor_sync(0xffffffffu, hi_y, o));
            cnt += __shfl_xor_sync(0xffffffffu, cnt, o);
        }
        if ((threadIdx.x & 31) == 0 && cnt > 0) {
            atomicMin(&t.ctl[kBhMinX], lo_x); atomicMax(&t.ctl[kBhMaxX], hi_x);
            atomicMin(&t.ctl[kBhMinY], lo_y); atomicMax(&t.ctl[kBhMaxY], hi_y);
            atomicAdd(&t.ctl[kBhLive], cnt);
        }
    }
    grid.sync();
    const int n_live = (int)t.ctl[kBhLive];
    if (n_live == 0) return;  // uniform: every thread reads the same word
    if (gid == 0) {
        // root square, SpatialPartition.hpp:50-56 (float arithmetic)
        const float min_x = float_from_order_key(t.ctl[kBhMinX]), max_x = float_from_order_key(t.ctl[kBhMaxX]);
        const float min_y = float_from_order_key(t.ctl[kBhMinY]), max_y = float_from_order_key(t.ctl[kBhMaxY]);
        float half = __fmul_rn(fmaxf(__fsub_rn(max_x, min_x), __fsub_rn(max_y, min_y)), 0.5f);
        if (half <= 0.0f) half = 1.0f;
        t.centre[0] = make_float2(__fmul_rn(__fadd_rn(min_x, max_x), 0.5f), __fmul_rn(__fadd_rn(min_y, max_y), 0.5f));
        t.cell[0] = make_float4(0.f, 0.f, 0.f, half);
        t.count[0] = n_live;
        t.level_begin[0] = 0;
        t.level_begin[1] = 1;
    }
    grid.sync();
    bh_stamp(t, gid, stamp);

    // ---- grow the tree one level per iteration (the fixed point of the insertions, :109-167)
    int lb = 0, le = 1, level = 0;
    for (;;) {
        // cells of this level holding two or more bodies get their four children (:158-167)
        for (long long k = lb + gid; k < le; k += gsz) {
            if (t.count[k] < 2) continue;
            const int base = (int)atomicAdd(&t.ctl[kBhNodes], 4u);
            if (base + 4 > t.capacity) {
                t.ctl[kBhError] = kBhOverflow;
                continue;
            }
            const float2 c = t.centre[k];
            const float hs = __fmul_rn(t.cell[k].w, 0.5f);
            const int skip = t.link[k].y;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                t.centre[base + q] = make_float2((q & 1) ? __fadd_rn(c.x, hs) : __fsub_rn(c.x, hs),
                                                 (q & 2) ? __fadd_rn(c.y, hs) : __fsub_rn(c.y, hs));
                t.cell[base + q] = make_float4(0.f, 0.f, 0.f, hs);
                t.count[base + q] = 0;
                // walk order is 3,2,1,0: child q is followed by child q-1, child 0 by whatever follows the parent
                t.link[base + q] = make_int2(-1, q > 0 ? base + q - 1 : skip);
            }
            t.link[k].x = base;
        }
        grid.sync();
        bh_stamp(t, gid, stamp);
        const int new_le = (int)min(t.ctl[kBhNodes], (unsigned int)t.capacity);
        const bool stop = new_le == le || t.ctl[kBhError] != kBhOk || level + 1 >= kBhMaxLevels;
        if (stop) {
            if (gid == 0) {
                t.ctl[kBhDepth] = (unsigned int)level;
                if (new_le != le && t.ctl[kBhError] == kBhOk) t.ctl[kBhError] = kBhTooDeep;
            }
            break;
        }
        // every body still travelling steps into the child that contains it (:127,:148,:153). Near the root a
        // level has few cells and every body is still travelling: counting straight into global memory would
        // serialise a million atomics on a handful of addresses (measured: 0.4-0.6 ms per level for the first
        // four levels at N = 2^20), so those levels count in a per-CTA shared histogram first.
        const int n_new = new_le - le;
        const bool hist = n_new <= kBhHistCells;
        if (hist) {
            for (int k = threadIdx.x; k < n_new; k += blockDim.x) s_hist[k] = 0;
            __syncthreads();
        }
        for (long long i = gid; i < n; i += gsz) {
            const int c = t.cur[i];
            if (c < 0) continue;
            const int first = t.link[c].x;
            if (first < 0) {  // alone in its cell: placed, and the cell learns whose it is
                t.link[c].x = -2 - (int)i;
                t.cur[i] = -1;
                if (level > 0 && level < kBhKeyLevels) t.key[i] <<= 2 * (kBhKeyLevels - level);
                continue;
            }
            const int q = bh_quadrant(t.centre[c], t.bpos[i]);
            const int child = first + q;
            t.cur[i] = child;
            if (hist) atomicAdd(&s_hist[child - le], 1);
            else atomicAdd(&t.count[child], 1);
            if (level < kBhKeyLevels) t.key[i] = (t.key[i] << 2) | (unsigned int)q;
        }
        if (hist) {
            __syncthreads();
            for (int k = threadIdx.x; k < n_new; k += blockDim.x)
                if (s_hist[k]) atomicAdd(&t.count[le + k], s_hist[k]);
        }
        lb = le;
        le = new_le;
        ++level;
        if (gid == 0) t.level_begin[level + 1] = le;
        grid.sync();
        bh_stamp(t, gid, stamp);
    }
    grid.sync();  // error word, level_begin and depth visible to everyone
    if (t.ctl[kBhError] != kBhOk) return;  // uniform
    // bodies still travelling sit alone in cells of the deepest level: place them (keys left-aligned like the others)
    for (long long i = gid; i < n; i += gsz) {
        const int c = t.cur[i];
        if (c < 0) continue;
        t.link[c].x = -2 - (int)i;
        if (level > 0 && level < kBhKeyLevels) t.key[i] <<= 2 * (kBhKeyLevels - level);
    }
    grid.sync();

    // ---- mass and centre of mass, deepest level first (:178-220)
    for (int lv = level; lv >= 0; --lv) {
        const int b = t.level_begin[lv], e = t.level_begin[lv + 1];
        for (long long k = b + gid; k < e; k += gsz) {
            const int first = t.link[k].x;
            const float half = t.cell[k].w;
            if (first < 0) {
                if (first == -1) {
                    t.cell[k] = make_float4(0.f, 0.f, 0.f, half);
                } else {
                    const int body = -2 - first;
                    const float2 p = t.bpos[body];
                    t.cell[k] = make_float4(p.x, p.y, s.mass[body], half);
                }
                continue;
            }
            float ms = 0.f, sx = 0.f, sy = 0.f;
#pragma unroll
            for (int q = 0; q < 4; ++q) {
                const float4 ch = t.cell[first + q];
                if (ch.z > 0.f) {
                    ms = __fadd_rn(ms, ch.z);
                    sx = __fadd_rn(sx, __fmul_rn(ch.x, ch.z));
                    sy = __fadd_rn(sy, __fmul_rn(ch.y, ch.z));
                }
            }
            float cx = 0.f, cy = 0.f;
            if (ms > 0.f) {
                const float inv = __fdiv_rn(1.0f, ms);
                cx = __fmul_rn(sx, inv);
                cy = __fmul_rn(sy, inv);
            }
            t.cell[k] = make_float4(cx, cy, ms, half);
        }
        if (lv > 0) grid.sync();
    }
    bh_stamp(t, gid, stamp);
    if (gid == 0 && t.trace) t.trace[0] = (unsigned long long)stamp;
}

// One walk per body (Physics.hpp:224-230 + SpatialPartition.hpp:63-106). order[]: body rows in Z-order.
// visits (optional): per-launch totals {cells popped, terms added, exact re-evaluations of the accept test}.
template <bool STATS>
__global__ void __launch_bounds__(kBhWalkThreads, NB_BH_WALK_MINB) bh_walk_kernel(StateView s, BhTree t, const int* __restrict__ order,
                                                                 double g, float softening, float theta_f,
                                                                 unsigned long long* visits) {
    const long long tid = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    unsigned long long n_pop = 0, n_term = 0, n_exact = 0;
    if (tid < s.n) {
        const int i = order[tid];
        if (s.live[i]) {
            float afx = 0.f, afy = 0.f;
            if (!s.pinned[i]) {  // Physics.hpp:225
                const float2 p = t.bpos[i];
                const double tx = (double)p.x, ty = (double)p.y;
                const double eps2 = __dmul_rn((double)softening, (double)softening);  // :186
                const double theta = (double)theta_f;                                   // :222
                const float th_lo = theta_f * (1.0f - 1e-5f), th_hi = theta_f * (1.0f + 1e-5f);
                const bool fast = t.cell[0].w < 1e30f;  // 2*half cannot overflow a float
                const float4* __restrict__ cell = t.cell;
                const int2* __restrict__ link = t.link;
                int node = 0;
                // One cell per trip. The trip is written as straight-line code with two rarely/partly taken
                // blocks (the FP64 re-evaluation of a borderline accept test, the accepted term) so that a warp
                // pays for two reconvergence points per cell instead of one per source-level branch.
                while (node >= 0) {
                    const float4 c = __ldg(&cell[node]);
                    const int2 l = __ldg(&link[node]);
                    if (STATS) ++n_pop;
                    const bool massive = c.z > 0.f;   // :72 (an empty leaf has mass 0)
                    const bool inner = l.x >= 0;
                    // accept test :88-91 in FP32 (harmless on a leaf: its result is not used there)
                    const float fdx = c.x - p.x, fdy = c.y - p.y;
                    const float d2 = fmaf(fdx, fdx, fdy * fdy);
                    float rs;
                    asm("rsqrt.approx.ftz.f32 %0, %1;" : "=f"(rs) : "f"(d2));
                    const float q = (c.w + c.w) * rs;
                    const bool decisive = fast && d2 > 1e-30f && d2 < 1e30f && (q < th_lo || q > th_hi);
                    bool accept = q < th_lo;
                    if (massive && inner && !decisive) {  // a handful of cells per million
                        const double dx = __dsub_rn((double)c.x, tx), dy = __dsub_rn((double)c.y, ty);
                        const double dist = __dsqrt_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)));
                        accept = __ddiv_rn(__dmul_rn((double)c.w, 2.0), dist) < theta;
                        if (STATS) ++n_exact;
                    }
                    // a leaf contributes unless it is empty or the target itself (:75); a cell when accepted
                    const bool term = massive && (inner ? accept : (-2 - l.x) != i);
                    node = (massive && inner && !accept) ? l.x + 3 : l.y;  // open: children 3,2,1,0
                    if (term) {
                        const double dx = __dsub_rn((double)c.x, tx), dy = __dsub_rn((double)c.y, ty);
                        const double r2 = __dadd_rn(__dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy)), eps2);
                        const double inv_r = __ddiv_rn(1.0, __dsqrt_rn(r2));
                        const double inv_r3 = __dmul_rn(__dmul_rn(inv_r, inv_r), inv_r);
                        const double gm = __dmul_rn(g, (double)c.z);
                        afx = __fadd_rn(afx, __double2float_rn(__dmul_rn(__dmul_rn(gm, dx), inv_r3)));
                        afy = __fadd_rn(afy, __double2float_rn(__dmul_rn(__dmul_rn(gm, dy), inv_r3)));
                        if (STATS) ++n_term;
                    }
                }
            }
            s.a[0][i] = (double)afx;  // :228-229, :260; a pinned body gets 0
            s.a[1][i] = (double)afy;
            s.a[2][i] = 0.0;
        }
    }
    if (STATS) {
#pragma unroll
        for (int o = 16; o > 0; o >>= 1) {
            n_pop += __shfl_xor_sync(0xffffffffu, n_pop, o);
            n_term += __shfl_xor_sync(0xffffffffu, n_term, o);
            n_exact += __shfl_xor_sync(0xffffffffu, n_exact, o);
        }
        if ((threadIdx.x & 31) == 0) {
            atomicAdd(&visits[0], n_pop);
            atomicAdd(&visits[1], n_term);
            atomicAdd(&visits[2], n_exact);
        }
    }
}

}  // namespace nb
