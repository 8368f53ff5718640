/* oracle/bh_oracle.c — TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C restatement of the Barnes-Hut branch of the reference's compute_gravity
 * (src/systems/Physics.hpp:212-230) and of the quadtree it uses
 * (src/physics/SpatialPartition.hpp:36-220). That branch is what the reference runs for every
 * world of more than bh_threshold (default 100, Constants.hpp:90) bodies, so it — not the direct
 * sum — is the reference's behaviour at its default settings for all but tiny worlds.
 *
 * Parity status: PINNED. tests/test_oracle.py compares the accelerations bit-for-bit with the
 * unmodified reference (oracle/_ref/libnbref.so) and with tests/golden/bh_*.npz.
 *
 * What has to be reproduced for bit parity:
 *   - positions and masses enter the tree as FLOATS (Physics.hpp:217): float(x), float(y);
 *   - the root square: centre ((min+max)*0.5f), half size max(dx,dy)*0.5f, 1 if that is <= 0
 *     (SpatialPartition.hpp:40-56), all in float;
 *   - a point-region quadtree with one body per leaf; a node that receives a second body gets all
 *     four children (:158-167), child k = (east ? 1 : 0) + (south ? 2 : 0) with east = x > cx,
 *     south = y > cy (:169-176), child centres c -+ half/2 in float. The shape depends only on the
 *     SET of points, not on insertion order, so any construction order gives the same tree;
 *   - float mass / centre-of-mass sums over the children in the order 0,1,2,3, skipping children
 *     with mass <= 0, com = sum * (1.0f / mass) (:201-212);
 *   - the walk (:63-106): depth-first with an explicit stack, children pushed 0..3 and therefore
 *     visited 3,2,1,0; a cell is accepted when (2*half)/|com - x| < theta with the distance
 *     UNsoftened in FP64; accepted cells and leaves add G*m*d/(r^2+eps^2)^(3/2) evaluated in FP64
 *     as ((G*m)*d)*invR3, each term ROUNDED TO FLOAT and accumulated in a float pair (:83-84,97-98);
 *   - pinned bodies are sources but not targets (:225).
 * Two bodies with identical float coordinates make the reference subdivide for ever (it has no
 * depth limit); this restatement stops at kMaxDepth and returns -1 instead.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

enum { kMaxDepth = 512 };

typedef struct {
    float cx, cy, half;   /* the square */
    float mass, comx, comy;
    int32_t body;         /* index into the gathered list, -1 if none */
    int32_t child;        /* first of four consecutive children, -1 for a leaf */
    int32_t depth;
} bh_node;

typedef struct {
    bh_node* v;
    int64_t len, cap;
} bh_arena;

static int64_t arena_push(bh_arena* a, float cx, float cy, float half, int depth) {
    if (a->len == a->cap) {
        a->cap = a->cap ? a->cap * 2 : 1024;
        a->v = (bh_node*)realloc(a->v, sizeof(bh_node) * (size_t)a->cap);
    }
    bh_node* nd = &a->v[a->len];
    nd->cx = cx; nd->cy = cy; nd->half = half;
    nd->mass = 0.0f; nd->comx = 0.0f; nd->comy = 0.0f;
    nd->body = -1; nd->child = -1; nd->depth = depth;
    return a->len++;
}

static int quadrant(const bh_node* nd, float x, float y) {
    /* SpatialPartition.hpp:169-176 */
    return (x > nd->cx ? 1 : 0) + (y > nd->cy ? 2 : 0);
}

static void split(bh_arena* a, int64_t k) {
    /* SpatialPartition.hpp:158-167: NW, NE, SW, SE */
    const float hs = a->v[k].half * 0.5f;
    const float cx = a->v[k].cx, cy = a->v[k].cy;
    const int d = a->v[k].depth + 1;
    const int64_t c0 = arena_push(a, cx - hs, cy - hs, hs, d);
    arena_push(a, cx + hs, cy - hs, hs, d);
    arena_push(a, cx - hs, cy + hs, hs, d);
    arena_push(a, cx + hs, cy + hs, hs, d);
    a->v[k].child = (int32_t)c0;
}

/* stats (optional, 4 values): nodes, deepest level, cells/leaves popped by all walks, terms added. */
int64_t orc_bh_gravity(int64_t n, const double* pos, const double* vel, const float* mass, const uint8_t* pinned,
                       double g, float softening, float theta_f, double* acc, int64_t* stats) {
    const double eps2 = (double)softening * (double)softening; /* Physics.hpp:186 */
    const double theta = (double)theta_f;                      /* :222 */
    /* gather, :197-205 and :213-218 */
    int64_t* idx = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    int64_t m = 0;
    for (int64_t i = 0; i < n; ++i) {
        const double* p = pos + 2 * i;
        const double* v = vel + 2 * i;
        if (isfinite(p[0]) && isfinite(p[1]) && isfinite(v[0]) && isfinite(v[1]) && mass[i] > 0.0f &&
            isfinite((double)mass[i]))
            idx[m++] = i;
    }
    if (stats) memset(stats, 0, 4 * sizeof(int64_t));
    if (m == 0) { free(idx); return 0; }
    float* bx = (float*)malloc(sizeof(float) * (size_t)m);
    float* by = (float*)malloc(sizeof(float) * (size_t)m);
    float* bm = (float*)malloc(sizeof(float) * (size_t)m);
    for (int64_t k = 0; k < m; ++k) {
        bx[k] = (float)pos[2 * idx[k]];
        by[k] = (float)pos[2 * idx[k] + 1];
        bm[k] = mass[idx[k]];
    }
    /* root square, SpatialPartition.hpp:40-56 */
    float min_x = bx[0], max_x = bx[0], min_y = by[0], max_y = by[0];
    for (int64_t k = 0; k < m; ++k) {
        if (bx[k] < min_x) min_x = bx[k];
        if (bx[k] > max_x) max_x = bx[k];
        if (by[k] < min_y) min_y = by[k];
        if (by[k] > max_y) max_y = by[k];
    }
    const float ex = max_x - min_x, ey = max_y - min_y;
    float half = (ex > ey ? ex : ey) * 0.5f;
    if (half <= 0.0f) half = 1.0f;
    bh_arena a = {0, 0, 0};
    arena_push(&a, (min_x + max_x) * 0.5f, (min_y + max_y) * 0.5f, half, 0);
    int64_t deepest = 0;
    int failed = 0;
    /* insertion, :109-156: walk down; an occupied leaf is split and its body moved one level down */
    for (int64_t k = 0; k < m && !failed; ++k) {
        int64_t cur = 0;
        for (;;) {
            bh_node* nd = &a.v[cur];
            if (nd->child < 0) {
                if (nd->body < 0) { nd->body = (int32_t)k; break; }
                if (nd->depth >= kMaxDepth) { failed = 1; break; }
                const int32_t old = nd->body;
                split(&a, cur);
                nd = &a.v[cur]; /* the arena may have moved */
                nd->body = -1;
                a.v[nd->child + quadrant(nd, bx[old], by[old])].body = old;
                if (nd->depth + 1 > deepest) deepest = nd->depth + 1;
            }
            cur = nd->child + quadrant(nd, bx[k], by[k]);
        }
    }
    if (failed) {
        free(a.v); free(bx); free(by); free(bm); free(idx);
        return -1;
    }
    /* mass and centre of mass, :178-220. Children always sit behind their parent in the arena, so
     * a backward sweep visits them first (the reference's post-order gives the same values). */
    for (int64_t k = a.len - 1; k >= 0; --k) {
        bh_node* nd = &a.v[k];
        if (nd->child < 0) {
            if (nd->body >= 0) { nd->mass = bm[nd->body]; nd->comx = bx[nd->body]; nd->comy = by[nd->body]; }
            continue;
        }
        float ms = 0.0f, sx = 0.0f, sy = 0.0f;
        for (int q = 0; q < 4; ++q) {
            const bh_node* ch = &a.v[nd->child + q];
            if (ch->mass > 0.0f) {
                ms += ch->mass;
                sx += ch->comx * ch->mass;
                sy += ch->comy * ch->mass;
            }
        }
        nd->mass = ms;
        if (ms > 0.0f) {
            const float inv = 1.0f / ms;
            nd->comx = sx * inv;
            nd->comy = sy * inv;
        }
    }
    /* the walks, Physics.hpp:224-230 + SpatialPartition.hpp:63-106 */
    int64_t* stack = (int64_t*)malloc(sizeof(int64_t) * (size_t)(3 * (deepest + 2) + 8));
    int64_t popped = 0, terms = 0;
    for (int64_t k = 0; k < m; ++k) {
        float afx = 0.0f, afy = 0.0f;
        if (!(pinned && pinned[idx[k]])) {
            const double tx = (double)bx[k], ty = (double)by[k];
            int64_t top = 0;
            stack[top++] = 0;
            while (top > 0) {
                const bh_node* nd = &a.v[stack[--top]];
                ++popped;
                if (nd->mass <= 0.0f) continue;
                const int leaf = nd->child < 0;
                if (leaf && (nd->body < 0 || nd->body == k)) continue;
                const double dx = (double)nd->comx - tx; /* a leaf's com IS its body's position */
                const double dy = (double)nd->comy - ty;
                if (!leaf) {
                    const double dist = sqrt(dx * dx + dy * dy);
                    if (!(((double)nd->half * 2.0) / dist < theta)) {
                        for (int q = 0; q < 4; ++q) stack[top++] = nd->child + q;
                        continue;
                    }
                }
                const double r2 = dx * dx + dy * dy + eps2;
                const double inv_r = 1.0 / sqrt(r2);
                const double inv_r3 = inv_r * inv_r * inv_r;
                const double gm = g * (double)nd->mass;
                afx += (float)(gm * dx * inv_r3);
                afy += (float)(gm * dy * inv_r3);
                ++terms;
            }
        }
        acc[2 * idx[k]] = (double)afx; /* :228-229, :260 */
        acc[2 * idx[k] + 1] = (double)afy;
    }
    if (stats) { stats[0] = a.len; stats[1] = deepest; stats[2] = popped; stats[3] = terms; }
    const int64_t n_nodes = a.len;
    free(stack); free(a.v); free(bx); free(by); free(bm); free(idx);
    return n_nodes;
}
