// common.cuh — shared device/host definitions for the B200 physics step.
//
// Data layout in HBM (all owned by nb_ctx, see nbody_b200.cu):
//   state   : SoA FP64, three components each: x[3][n], v[3][n], a[3][n], a0[3][n]
//             (the reference's Position/Velocity/Acceleration/PrevAcceleration DVec2
//              components, src/components/Components.hpp:10-21, widened with z)
//   mass    : f32[n] (Components.hpp:22-24), pinned : u8[n] (Components.hpp:36-38)
//   j-stream: FP32 "pair-SoA" packed sources, 32 B per body PAIR:
//             {x0,x1,y0,y1,z0,z1,m0,m1}; two generations (cur/next) because the fused
//             epilogue writes step k+1's stream while other CTAs still read step k's.
//             Positions are stored RELATIVE TO THE FP64 ORIGIN OF THEIR GROUP of kGroupBodies
//             consecutive stream rows (GroupHdr), scaled by an exact power of two 2^-e, as a
//             float `hi` (pk) plus a float residual `lo` (pk_lo): 48 bits of position relative
//             to the group's extent, whatever the distance of the group from the world origin.
//             Masses are scaled by 2^-q; G, 2^q and 2^-2e are applied in FP64 in the epilogue.
//   order   : stream row -> body row. The stream is in Hilbert-curve order of the positions so a
//             group is spatially compact; the FP64 state stays in upload order.
#pragma once
#include <cuda_runtime.h>
#include <stdint.h>

namespace nb {

#ifndef NB_TILE
#define NB_TILE 512
#endif
constexpr int kTileBodies = NB_TILE;      // j-bodies per shared-memory stage (16 B each: 8 KB)
#ifndef NB_STAGES
#define NB_STAGES 2
#endif
#ifndef NB_RING_SYNC
#define NB_RING_SYNC 1  // 1: hand stages back with a CTA-wide barrier instead of per-warp `empty` mbarriers
#endif
constexpr int kStages = NB_STAGES;         // TMA bulk-copy ring depth (8 KB stages)
constexpr int kFlushChunk = 64;           // FP32 partial sums are folded into FP64 every 64 sources
constexpr int kExtentSlots = 4;           // rotating max|x| slots, see scale_for()
constexpr int kHeadroomLog2 = 17;         // scaled |x| < 2^17 when packed; r^6 stays finite up to 2^19
constexpr int kMaxSegments = 64;          // j-split factor upper bound
constexpr float kFarPacked = 1152921504606846976.0f;  // 2^60: packed x of non-attracting rows
constexpr int kDiagValues = 12;           // KE, PE, Px,Py,Pz, Cx,Cy,Cz, M, nonfinite, pad, pad
constexpr int kGroupBodies = 128;         // stream rows that share one FP64 origin
constexpr int kGroupsPerTile = kTileBodies / kGroupBodies;
// A target closer to a source group's bounding box than kNearFactor x (the box's largest half-extent)
// takes the two-float (hi + lo) path for that group; everything farther is evaluated from `hi` alone with
// a relative error of dx below 2^-25 (1 + 2 / kNearFactor).
#ifndef NB_NEAR_FACTOR
#define NB_NEAR_FACTOR 0.25f
#endif
constexpr float kNearFactor = NB_NEAR_FACTOR;

// Origin and half-extents (packed units) of one group of the j-stream. 48 B: three per 16-byte line.
struct alignas(16) GroupHdr {
    float h[3];    // max |hi| per axis over the group's attracting rows, packed units
    float reach;   // kNearFactor * max(h)
    double c[3];   // FP64 origin in packed units (physical origin * 2^-e, exact)
    float pad[2];
};
static_assert(sizeof(GroupHdr) == 48, "GroupHdr must stay 48 bytes (bulk-copied with the tile)");
constexpr uint32_t kHdrBytesPerTile = kGroupsPerTile * sizeof(GroupHdr);

// Power-of-two scales for one generation of the j-stream.
struct Scale {
    double inv_len;   // 2^-e : physical length -> packed units
    double acc_unit;  // G * 2^q * 2^-2e : packed acceleration sum -> m/s^2
    double pot_unit;  // 2^q * 2^-e : packed potential sum -> kg/m
    float eps2;       // softening^2 in packed units, floored so r^6 never underflows
};

__host__ __device__ inline double pow2d(int k) {
    // exact 2^k for k in the normal double range
    union { uint64_t u; double d; } c;
    c.u = (uint64_t)(k + 1023) << 52;
    return c.d;
}

// extent_bits: IEEE bits of max|coordinate| over live bodies (a non-negative float).
__host__ __device__ inline int extent_exponent(unsigned int extent_bits) {
    int ex = (int)((extent_bits >> 23) & 0xffu) - 126;  // value < 2^ex for normal floats
    if (ex < -60) ex = -60;                            // all bodies at (or next to) the origin
    if (ex > 100) ex = 100;
    return ex;
}

__host__ __device__ inline Scale scale_for(unsigned int extent_bits, int mass_exp, double g, float softening) {
    const int e = extent_exponent(extent_bits) - kHeadroomLog2;  // length unit 2^e
    Scale s;
    s.inv_len = pow2d(-e);
    s.acc_unit = g * pow2d(mass_exp) * s.inv_len * s.inv_len;
    s.pot_unit = pow2d(mass_exp) * s.inv_len;
    const double eps = (double)softening * s.inv_len;
    double eps2 = eps * eps;
    const double floor2 = 4.547473508864641e-13;  // 2^-41: (2^-41)^3 = 2^-123 > FLT_MIN
    if (!(eps2 >= floor2)) eps2 = floor2;
    if (eps2 > 1.0e30) eps2 = 1.0e30;
    s.eps2 = (float)eps2;
    return s;
}

// Where a pack writes one generation of the j-stream: the device's own buffers and, in a single-process
// multi-GPU context (nb_create_multi), the same buffers of every peer device — the all-gather of the packed
// positions is done by the producing kernel's own stores over NVLink (peer memory), not by a collective.
// extent[]: each device's copy of the rotating max|x| slots (max-reduced with peer atomics the same way).
constexpr int kMaxPeers = 8;
struct PeerSet {
    int n;
    float* pk[kMaxPeers];
    float* lo[kMaxPeers];
    GroupHdr* hdr[kMaxPeers];
    unsigned int* extent[kMaxPeers];
};

// Epilogue modes of the force kernel.
enum Mode : int {
    kForceOnly = 0,     // Physics::compute_gravity, Physics.hpp:183-261
    kVerletKick = 1,    // ... + kick of Physics.hpp:309-323
    kVerletKickDrift = 2,  // ... + drift of the NEXT substep (Physics.hpp:297-304) and its pack
    kEuler = 3,         // ... + kick/drift of Physics.hpp:276-290 and the next pack
    kPartialOnly = 4,   // no epilogue: leave the FP64 sums in `part` slot 0 for a later launch
    kDiagOnly = 5       // potential-only kernels: per-body diagnostics terms, nothing written to the state
};

struct StepParams {
    double* x[3];               // positions of generation `gen` (what the force pass evaluates)
    double* xs[3];              // kVerletKickDrift: the NEXT substep's drifted positions go here (speculative:
                                // committed by swapping x/xs when the next substep really uses the same dt)
    double* v[3];
    double* a[3];
    double* a0[3];
    const float* mass;
    const uint8_t* pinned;
    const uint8_t* live;        // gather filter of Physics.hpp:198-199 for generation `gen`, per body row
    uint8_t* live_next;         // ... for generation gen+1 (written together with the pack)
    const int* order;           // stream row -> body row, rows [i_lo, i_hi)
    const double* tx_cur[3];    // FP64 positions of this rank's rows IN STREAM ORDER (generation `gen`): the targets' coordinates,
                                // read coalesced by every (tile, segment) CTA instead of gathered through `order`
    double* tx_next[3];         // ... of generation gen+1, written by the pack
    const float* pk_cur;        // pair-SoA j-stream being read (generation `gen`): hi parts
    const float* lo_cur;        // pair-SoA residuals {xl0,xl1,yl0,yl1,zl0,zl1,0,0}, read by the near path only
    const GroupHdr* hdr_cur;    // one per kGroupBodies stream rows
    PeerSet out;                // generation gen+1, written by the epilogue (on every device of the context)
    const unsigned int* extent; // kExtentSlots rotating slots of float bits (this device's copy; written through `out`)
    double* part;               // [1 + nseg][4][i_cap] FP64 partial sums: slot 0 carries sums between two
                                // launches (local / remote sources), slots 1.. the source segments
    unsigned int* tile_counter; // [n_itiles] arrival counters (only nseg > 1)
    double* diag_part;          // [n_itiles][kDiagValues] per-CTA diagnostics partials
    int* filter_flip;           // pinned host word (or null): set when a kick changes a body's gather-filter outcome
    unsigned long long* stats;  // null, or stats[1] counts the (warp, group) visits that took the near path
    long long n;                // bodies
    long long n_tiles;          // j-stream length in tiles of kTileBodies (padded with m = 0)
    long long i_lo, i_hi;       // target range owned by this rank
    long long i_cap;            // row stride of `part`
    long long j_tile_lo, j_tile_hi;  // LOGICAL source tiles this launch covers, split over nseg CTAs
    long long hole_lo, hole_len;     // logical tile t >= hole_lo maps to physical tile t + hole_len
                                     // (the remote-sources launch skips the rank's own tiles this way)
    long long row_cap;          // end of this rank's padded stream rows (groups at or beyond it are never packed)
    int nseg;
    int gen;
    int slot_cur, slot_nxt;     // extent slots: scale of the stream being read / of the one the epilogue packs
    int mode;
    int accumulate;             // 1: add `part` slot 0 (an earlier launch's sums) before the epilogue
    int mass_exp;
    float softening;
    float max_speed;
    double g;
    double dt;                  // (double)dtSub
    double half_dt2;            // 0.5 * dt * dt, Physics.hpp:300
    double eqm_mass;            // uniform-mass kernels: the common packed mass (1.0 otherwise)
};

}  // namespace nb
