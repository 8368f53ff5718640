// merge.cuh — the merges of the reference's collision pass (Collision::resolve, merge mode,
// src/systems/Collision.hpp:75-132) replayed ON THE DEVICE in the reference's visiting order.
//
// The reference walks all pairs (i, j > i) in lexicographic order and merges the ones that overlap as
// it goes; a merge changes the survivor's position, mass and radius, so the outcome of later pair
// tests depends on the earlier merges (and a row keeps running with its body's snapshot after that
// body died, :76). The replay is an event queue over exactly the pairs that can overlap:
//   * the pairs that overlap in the initial state (detected by the kernels of collide.cuh), sorted by
//     key i*n + j (cub radix sort): consumed front to back;
//   * after every merge, the pairs (survivor, k) whose current states overlap and whose key lies AHEAD of
//     the event being processed: found by a neighbourhood scan of the uniform grid of the broad phase
//     (or of all bodies when the world has no grid) and pushed on a binary min-heap.
// Events are popped in key order and re-tested with the CURRENT snapshots; a pair that overlaps when the
// reference would test it was detected either initially or by the scan after the last change of one of
// its bodies, so the replay performs exactly the reference's merges, in its order (merge_log), with its
// arithmetic: every FP64 operation is the non-contracted one of the reference's -O2 build, and
// update_radius_from_mass' std::cbrt (:37-46) is glibc 2.39's algorithm restated operation by operation
// (cbrt_glibc below; pinned against libm by tests/test_collision_merge.py).
//
// One persistent CTA: thread 0 runs the sequential logic (a handful of L2-resident loads and ~60 FP64
// operations per event), all threads scan the survivor's neighbourhood. No host round trip per merge.
#pragma once
#include "collide.cuh"

namespace nb {

constexpr int kMergeThreads = 1024;

// glibc 2.39 sysdeps/ieee754/dbl-64/s_cbrt.c, restated with separately rounded operations.
__device__ __forceinline__ double cbrt_glibc(double x) {
    const double ax = fabs(x);
    if (!(ax > 0.0) || !(ax <= 1.7976931348623157e308)) return __dadd_rn(x, x);  // 0, inf, nan
    // frexp: ax = xm * 2^xe, xm in [0.5, 1)
    long long bits = __double_as_longlong(ax);
    int xe = (int)((bits >> 52) & 0x7ff);
    if (xe == 0) {  // subnormal: normalise first
        bits = __double_as_longlong(__dmul_rn(ax, 18014398509481984.0));  // 2^54
        xe = (int)((bits >> 52) & 0x7ff) - 54;
    }
    xe -= 1022;
    const double xm = __longlong_as_double((bits & 0x800fffffffffffffll) | 0x3fe0000000000000ll);
    double u = __dsub_rn(0.784932344976639262, __dmul_rn(0.145263899385486377, xm));
    u = __dsub_rn(1.83469277483613086, __dmul_rn(u, xm));
    u = __dsub_rn(2.44693122563534430, __dmul_rn(u, xm));
    u = __dsub_rn(2.11499494167371287, __dmul_rn(u, xm));
    u = __dsub_rn(1.50819193781584896, __dmul_rn(u, xm));
    u = __dadd_rn(0.354895765043919860, __dmul_rn(u, xm));
    const double t2 = __dmul_rn(__dmul_rn(u, u), u);
    const double f = (xe % 3 == -2)   ? 1.0 / 1.5874010519681994748
                     : (xe % 3 == -1) ? 1.0 / 1.2599210498948731648
                     : (xe % 3 == 0)  ? 1.0
                     : (xe % 3 == 1)  ? 1.2599210498948731648
                                      : 1.5874010519681994748;
    double ym = __dmul_rn(u, __dadd_rn(t2, __dmul_rn(2.0, xm)));
    ym = __ddiv_rn(ym, __dadd_rn(__dmul_rn(2.0, t2), xm));
    ym = __dmul_rn(ym, f);
    // ldexp(ym, xe / 3): exact for the normal results this code sees
    return scalbn(x > 0.0 ? ym : -ym, xe / 3);
}

__global__ void cbrt_probe_kernel(const double* in, double* out, int n) {
    const int i = blockIdx.x * blockDim.x + threadIdx.x;
    if (i < n) out[i] = cbrt_glibc(in[i]);
}

struct MergeParams {
    long long n;
    // current snapshots (Collision.hpp:51-67), updated by the merges
    double* cx; double* cy; double* vx; double* vy; double* cr;
    float* cm;
    unsigned char* pin;
    unsigned char* alive;    // 1 until the body is absorbed (:124-125)
    unsigned char* live;     // passes the input filter and is alive: what a scan may hit
    unsigned char* changed;
    const double* density;
    const double* x_init; const double* y_init;  // the positions the grid was binned with
    // events
    const unsigned long long* initial;  // sorted keys of the initially overlapping pairs
    unsigned long long n_initial;
    unsigned long long* heap;
    unsigned long long heap_cap;
    unsigned int* hits;                 // scratch of one scan, n entries
    // grid of the broad phase (cells[0] == 0: none, scan all bodies)
    GridSpec g;
    const unsigned int* cell_start;
    const unsigned int* sorted_idx;
    double r_max;                       // largest radius at the start
    // results
    long long* merge_log;               // {survivor, absorbed} per merge, in order
    long long merge_log_cap;
    long long* out;                     // [0] merges, [1] heap overflow flag, [2] events processed, [3] scan candidates
};

__device__ __forceinline__ void heap_push(unsigned long long* h, unsigned long long& size, unsigned long long key) {
    unsigned long long i = size++;
    while (i > 0) {
        const unsigned long long parent = (i - 1) >> 1;
        const unsigned long long pv = h[parent];
        if (pv <= key) break;
        h[i] = pv;
        i = parent;
    }
    h[i] = key;
}
__device__ __forceinline__ unsigned long long heap_pop(unsigned long long* h, unsigned long long& size) {
    const unsigned long long top = h[0];
    const unsigned long long last = h[--size];
    unsigned long long i = 0;
    for (;;) {
        unsigned long long c = 2 * i + 1;
        if (c >= size) break;
        if (c + 1 < size && h[c + 1] < h[c]) ++c;
        if (h[c] >= last) break;
        h[i] = h[c];
        i = c;
    }
    if (size > 0) h[i] = last;
    return top;
}

__global__ void __launch_bounds__(kMergeThreads, 1) merge_replay_kernel(const MergeParams p) {
    __shared__ long long s_sv;            // survivor of the merge just applied, or -1: no scan needed
    __shared__ double s_x, s_y, s_r, s_reach;
    __shared__ unsigned int s_nhits;
    __shared__ int s_done;
    const int tid = threadIdx.x;
    const unsigned long long un = (unsigned long long)p.n;
    const double pi = 3.141592653589793238462643383279502884;  // std::numbers::pi

    // thread 0's sequential state
    unsigned long long ptr = 0, heap_size = 0, key = 0, last_key = ~0ull;
    long long cur_row = -1, merges = 0, events = 0, candidates = 0;
    bool row_alive = false, overflow = false;
    double r_max_now = p.r_max, max_shift = 0.0;

    for (;;) {
        if (tid == 0) {
            s_sv = -1;
            s_done = 0;
            // ---- pop events until one merges (or none is left) ----------------------------------
            for (;;) {
                const bool have_a = ptr < p.n_initial, have_h = heap_size > 0;
                if (!have_a && !have_h) { s_done = 1; break; }
                const unsigned long long ka = have_a ? p.initial[ptr] : ~0ull;
                const unsigned long long kh = have_h ? p.heap[0] : ~0ull;
                if (ka <= kh) { key = ka; ++ptr; } else { key = heap_pop(p.heap, heap_size); }
                if (key == last_key) continue;  // the same pair reported by two scans: one event (std::set semantics)
                last_key = key;
                ++events;
                const long long a = (long long)(key / un), b = (long long)(key % un);
                if (a != cur_row) {  // alive[i] is looked at when row i starts and not again (:76)
                    cur_row = a;
                    row_alive = p.alive[a] != 0;
                }
                if (!row_alive || !p.alive[b]) continue;  // :76, :78
                const double xa = p.cx[a], ya = p.cy[a], xb = p.cx[b], yb = p.cy[b];
                const double dx = __dsub_rn(xb, xa), dy = __dsub_rn(yb, ya);
                const double rsum = __dadd_rn(p.cr[a], p.cr[b]);
                const double dist2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
                if (dist2 > __dmul_rn(rsum, rsum)) continue;  // :85, with the snapshots as they are NOW
                const float fa = p.cm[a], fb = p.cm[b];
                const bool pin_a = p.pin[a] != 0, pin_b = p.pin[b] != 0;
                const bool a_survives = (fa >= fb) || pin_b;  // :94
                const long long sv = a_survives ? a : b, dd = a_survives ? b : a;
                const bool pin_s = a_survives ? pin_a : pin_b, pin_d = a_survives ? pin_b : pin_a;
                if (pin_d && pin_s) continue;  // :97-104
                const double ms = (double)(a_survives ? fa : fb), md = (double)(a_survives ? fb : fa);
                const double xs = a_survives ? xa : xb, ys = a_survives ? ya : yb;
                const double xd = a_survives ? xb : xa, yd = a_survives ? yb : ya;
                const double M = __dadd_rn(ms, md);                                                     // :107
                const double px = __dadd_rn(__dmul_rn(ms, p.vx[sv]), __dmul_rn(md, p.vx[dd]));          // :108-109
                const double py = __dadd_rn(__dmul_rn(ms, p.vy[sv]), __dmul_rn(md, p.vy[dd]));
                const double xx = __ddiv_rn(__dadd_rn(__dmul_rn(ms, xs), __dmul_rn(md, xd)), M);        // :111-112
                const double xy = __ddiv_rn(__dadd_rn(__dmul_rn(ms, ys), __dmul_rn(md, yd)), M);
                const float mf = (float)M;                                                              // :115,128
                p.cm[sv] = mf;
                p.vx[sv] = __ddiv_rn(px, M);                                                            // :110,116
                p.vy[sv] = __ddiv_rn(py, M);
                p.cx[sv] = xx;
                p.cy[sv] = xy;
                p.pin[sv] = (pin_s || pin_d) ? 1 : 0;                                                   // :120,131
                const double rs = cbrt_glibc(__ddiv_rn(__dmul_rn(3.0, fmax(1.0, (double)mf)),
                                                      __dmul_rn(4.0 * pi, p.density[sv])));              // :37-46,132
                p.cr[sv] = rs;
                p.changed[sv] = 1;
                p.alive[dd] = 0;                                                                        // :124-125
                p.live[dd] = 0;
                if (merges < p.merge_log_cap) {
                    p.merge_log[2 * merges] = sv;
                    p.merge_log[2 * merges + 1] = dd;
                }
                ++merges;
                // what the scan must cover: every living k whose CURRENT state can overlap the survivor's new
                // one. The grid knows k's initial cell; k has moved by at most max_shift since (only survivors move).
                r_max_now = fmax(r_max_now, rs);
                max_shift = fmax(max_shift, fmax(fabs(xx - p.x_init[sv]), fabs(xy - p.y_init[sv])));
                s_sv = sv;
                s_x = xx; s_y = xy; s_r = rs;
                s_reach = (rs + r_max_now + max_shift) * (1.0 + 1e-9);
                s_nhits = 0;
                break;
            }
        }
        __syncthreads();
        if (s_done) break;
        // ---- scan the survivor's neighbourhood (all threads) -----------------------------------------
        const long long sv = s_sv;
        const double sx = s_x, sy = s_y, sr = s_r;
        auto test = [&](long long k) {
            if (k == sv || !p.live[k]) return;
            // (x_B - x_A)^2 and r_A + r_B do not depend on which of the two is A: one form serves both orders
            const double dx = __dsub_rn(p.cx[k], sx);
            const double dy = __dsub_rn(p.cy[k], sy);
            const double rsum = __dadd_rn(sr, p.cr[k]);
            const double dist2 = __dadd_rn(__dmul_rn(dx, dx), __dmul_rn(dy, dy));
            if (!(dist2 > __dmul_rn(rsum, rsum))) p.hits[atomicAdd(&s_nhits, 1u)] = (unsigned int)k;
        };
        long long scanned = 0;
        if (p.g.cells[0] > 0) {
            const double reach = s_reach;
            const int x_lo = grid_axis_cell(sx - reach, p.g.origin[0], p.g.inv_h, p.g.cells[0]);
            const int x_hi = grid_axis_cell(sx + reach, p.g.origin[0], p.g.inv_h, p.g.cells[0]);
            const int y_lo = grid_axis_cell(sy - reach, p.g.origin[1], p.g.inv_h, p.g.cells[1]);
            const int y_hi = grid_axis_cell(sy + reach, p.g.origin[1], p.g.inv_h, p.g.cells[1]);
            for (int qy = y_lo; qy <= y_hi; ++qy) {
                const long long row = (long long)qy * p.g.cells[0];
                const unsigned int k_lo = p.cell_start[row + x_lo], k_hi = p.cell_start[row + x_hi + 1];
                for (unsigned int q = k_lo + tid; q < k_hi; q += kMergeThreads) test((long long)p.sorted_idx[q]);
                scanned += k_hi - k_lo;
            }
        } else {
            for (long long k = tid; k < p.n; k += kMergeThreads) test(k);
            scanned = p.n;
        }
        __syncthreads();
        // ---- queue the hits that still lie ahead in the visiting order (thread 0) ----------------------
        if (tid == 0) {
            candidates += scanned;
            const unsigned int nh = s_nhits;
            for (unsigned int h = 0; h < nh; ++h) {
                const unsigned long long k = p.hits[h];
                const unsigned long long lo = k < (unsigned long long)sv ? k : (unsigned long long)sv;
                const unsigned long long hi = k < (unsigned long long)sv ? (unsigned long long)sv : k;
                const unsigned long long nk = lo * un + hi;
                if (nk > key) {  // only tests the reference has not made yet
                    if (heap_size < p.heap_cap) heap_push(p.heap, heap_size, nk);
                    else overflow = true;
                }
            }
        }
        // no barrier needed here: the next iteration's thread-0 phase writes s_* only after it, and the other
        // threads wait at the barrier above before reading them
    }
    if (tid == 0) {
        p.out[0] = merges;
        p.out[1] = overflow ? 1 : 0;
        p.out[2] = events;
        p.out[3] = candidates;
    }
}

// scratch copies for the merge phase from the state and the caller's arrays
__global__ void __launch_bounds__(256) merge_snapshot_kernel(const double* __restrict__ x0, const double* __restrict__ x1,
                                                             const double* __restrict__ x2, const double* __restrict__ v0,
                                                             const double* __restrict__ v1, const float* __restrict__ mass,
                                                             const unsigned char* __restrict__ pinned, long long n,
                                                             MergeParams p) {
    const long long i = blockIdx.x * (long long)blockDim.x + threadIdx.x;
    if (i >= n) return;
    p.cx[i] = x0[i]; p.cy[i] = x1[i];
    p.vx[i] = v0[i]; p.vy[i] = v1[i];
    p.cm[i] = mass[i];
    p.pin[i] = pinned[i];
    p.alive[i] = 1;
    p.changed[i] = 0;
    p.live[i] = collide_input_ok(x0[i], x1[i], x2[i], mass[i]) ? 1 : 0;
}

}  // namespace nb
