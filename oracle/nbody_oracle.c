/* oracle/nbody_oracle.c — TEST INFRASTRUCTURE, not product code.
 *
 * Plain-C, FP64, single-threaded restatement of the reference's physics step on flat
 * arrays (no flecs). Only tests/, __graft_entry__.smoke() and bench.py's cpu_baseline /
 * --impl reference legs may call it, and only as the checker or the timed CPU baseline.
 * The product path (raylib-nbody_b200/csrc) never links or loads this file.
 *
 * Parity status: PINNED. tests/test_oracle.py checks every function here bit-for-bit
 * against the unmodified reference compiled into oracle/_ref/libnbref.so (in the build
 * container) and against fixtures that library generated (tests/golden/, everywhere).
 *
 * Array conventions follow the reference's component layouts:
 *   pos/vel/acc/acc_prev : n x {x,y} doubles  (DVec2, src/core/Math.hpp:6-9)
 *   mass                 : n floats           (Mass,  src/components/Components.hpp:22-24)
 *   pinned               : n bytes            (Pinned, Components.hpp:36-38)
 * The *3 variants carry {x,y,z}; with z == 0 they reduce exactly to the 2-D code
 * (every z term is +0.0), which is how the 3-D CUDA path is tied to the 2-D reference.
 */
#include <math.h>
#include <stdint.h>
#include <stdlib.h>
#include <string.h>

typedef struct {
    double g;                   /* Config.hpp:13 */
    float softening;            /* Config.hpp:15 */
    float max_speed;            /* Config.hpp:16, 0 = uncapped */
    int integrator;             /* Config.hpp:25, 0 = semi-implicit Euler, 1 = velocity Verlet */
    float max_substep;          /* Config.hpp:28 */
    int max_substeps_per_frame; /* Config.hpp:29 */
    int bh_threshold;           /* Config.hpp:17: Barnes-Hut above this many gathered bodies (2-D only) */
    float bh_theta;             /* Config.hpp:18 */
} orc_config;

static int body_is_live(const double* p, const double* v, float m, int dim) {
    /* gather filter of compute_gravity, Physics.hpp:198-199 */
    for (int k = 0; k < dim; ++k)
        if (!isfinite(p[k]) || !isfinite(v[k])) return 0;
    return m > 0.0f && isfinite((double)m);
}

/* Direct all-pairs softened gravity, Physics.hpp:183-261 (direct branch :231-258).
 * Bodies failing the gather filter neither attract nor get a new acceleration (their
 * acc entry is left untouched, as the reference never scatters to them, :260). */
static void gravity_dim(int64_t n, const double* pos, const double* vel, const float* mass,
                        const uint8_t* pinned, double g, float softening, double* acc, int dim) {
    const double eps2 = (double)softening * (double)softening; /* :186 */
    int64_t* idx = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    int64_t m = 0;
    for (int64_t i = 0; i < n; ++i)
        if (body_is_live(pos + dim * i, vel + dim * i, mass[i], dim)) idx[m++] = i;
    double* a = (double*)calloc((size_t)(m > 0 ? m : 1) * 3, sizeof(double));
    for (int64_t ii = 0; ii < m; ++ii) {
        const int64_t i = idx[ii];
        const double* pi = pos + dim * i;
        const int pin_i = pinned && pinned[i];
        for (int64_t jj = ii + 1; jj < m; ++jj) {
            const int64_t j = idx[jj];
            const double* pj = pos + dim * j;
            double d[3] = {0.0, 0.0, 0.0};
            for (int k = 0; k < dim; ++k) d[k] = pj[k] - pi[k];
            /* :239  r2 = dx*dx + dy*dy + eps2 ; the 3-D extension inserts dz*dz before eps2 */
            double r2 = d[0] * d[0] + d[1] * d[1];
            if (dim == 3) r2 += d[2] * d[2];
            r2 += eps2;
            const double inv_r = 1.0 / sqrt(r2);           /* :240 */
            const double inv_r3 = inv_r * inv_r * inv_r;   /* :241 */
            const double gmj = g * (double)mass[j];        /* :243  ((G*m)*dx)*invR3 */
            const double ngmi = -g * (double)mass[i];      /* :245 */
            if (!pin_i)
                for (int k = 0; k < dim; ++k) a[3 * ii + k] += gmj * d[k] * inv_r3;
            if (!(pinned && pinned[j]))
                for (int k = 0; k < dim; ++k) a[3 * jj + k] += ngmi * d[k] * inv_r3;
        }
    }
    for (int64_t ii = 0; ii < m; ++ii)
        for (int k = 0; k < dim; ++k) acc[dim * idx[ii] + k] = a[3 * ii + k];
    free(a);
    free(idx);
}

void orc_compute_gravity(int64_t n, const double* pos, const double* vel, const float* mass,
                         const uint8_t* pinned, double g, float softening, double* acc) {
    gravity_dim(n, pos, vel, mass, pinned, g, softening, acc, 2);
}

void orc_compute_gravity3(int64_t n, const double* pos, const double* vel, const float* mass,
                          const uint8_t* pinned, double g, float softening, double* acc) {
    gravity_dim(n, pos, vel, mass, pinned, g, softening, acc, 3);
}

/* Barnes-Hut branch of compute_gravity (oracle/bh_oracle.c). */
int64_t orc_bh_gravity(int64_t n, const double* pos, const double* vel, const float* mass, const uint8_t* pinned,
                       double g, float softening, float theta, double* acc, int64_t* stats);

/* The switch of Physics.hpp:212: the quadtree when more than bh_threshold bodies pass the gather
 * filter, the direct sum otherwise. The 3-D extension has no tree (the reference is 2-D). */
static void gravity_switch(int64_t n, const double* pos, const double* vel, const float* mass,
                           const uint8_t* pinned, const orc_config* cfg, double* acc, int dim) {
    if (dim == 2) {
        int64_t live = 0;
        for (int64_t i = 0; i < n; ++i) live += body_is_live(pos + 2 * i, vel + 2 * i, mass[i], 2);
        if (live > (int64_t)cfg->bh_threshold) {
            orc_bh_gravity(n, pos, vel, mass, pinned, cfg->g, cfg->softening, cfg->bh_theta, acc, NULL);
            return;
        }
    }
    gravity_dim(n, pos, vel, mass, pinned, cfg->g, cfg->softening, acc, dim);
}

void orc_compute_gravity_cfg(int64_t n, const double* pos, const double* vel, const float* mass,
                             const uint8_t* pinned, const orc_config* cfg, double* acc) {
    gravity_switch(n, pos, vel, mass, pinned, cfg, acc, 2);
}

/* Substep plan, Physics.hpp:268-271 (float arithmetic throughout). */
int orc_substep_plan(float dt, float max_substep, int max_substeps_per_frame, float* dt_sub) {
    const float cap = fmaxf(1e-6f, max_substep);
    int n_steps = (int)ceilf(dt / cap);
    const int hi = max_substeps_per_frame > 1 ? max_substeps_per_frame : 1;
    if (n_steps > hi) n_steps = hi;
    if (n_steps < 1) n_steps = 1;
    *dt_sub = dt / (float)n_steps;
    return n_steps;
}

static void cap_speed(double* v, float max_speed, int dim) {
    /* Physics.hpp:280-287 / :315-322 */
    if (max_speed > 0.0f) {
        double s2 = v[0] * v[0] + v[1] * v[1];
        if (dim == 3) s2 += v[2] * v[2];
        const double vlen = sqrt(s2);
        if (vlen > (double)max_speed) {
            const double s = (double)max_speed / vlen;
            for (int k = 0; k < dim; ++k) v[k] *= s;
        }
    }
}

/* integrate(), Physics.hpp:263-327. acc must hold a(x) on entry (sys2 ran before sys3). */
static void integrate_dim(int64_t n, double* pos, double* vel, double* acc, double* acc_prev,
                          const float* mass, const uint8_t* pinned, const orc_config* cfg, float dt,
                          int dim) {
    float dt_sub;
    const int n_steps = orc_substep_plan(dt, cfg->max_substep, cfg->max_substeps_per_frame, &dt_sub);
    const double h = (double)dt_sub; /* float dtSub promoted in every product below */
    if (cfg->integrator == 0) {
        for (int step = 0; step < n_steps; ++step) {
            for (int64_t i = 0; i < n; ++i) {
                if (pinned && pinned[i]) continue;
                for (int k = 0; k < dim; ++k) vel[dim * i + k] += acc[dim * i + k] * h; /* :278 */
                cap_speed(vel + dim * i, cfg->max_speed, dim);
                for (int k = 0; k < dim; ++k) pos[dim * i + k] += vel[dim * i + k] * h; /* :288 */
            }
            if (step + 1 < n_steps) /* :292 */
                gravity_switch(n, pos, vel, mass, pinned, cfg, acc, dim);
        }
    } else {
        const double half_dt2 = 0.5 * h * h; /* :300 */
        for (int step = 0; step < n_steps; ++step) {
            for (int64_t i = 0; i < n; ++i) {
                if (pinned && pinned[i]) continue;
                for (int k = 0; k < dim; ++k) {
                    pos[dim * i + k] += vel[dim * i + k] * h + acc[dim * i + k] * half_dt2; /* :301 */
                    acc_prev[dim * i + k] = acc[dim * i + k];                               /* :303 */
                }
            }
            gravity_switch(n, pos, vel, mass, pinned, cfg, acc, dim); /* :307 */
            for (int64_t i = 0; i < n; ++i) {
                if (pinned && pinned[i]) continue;
                for (int k = 0; k < dim; ++k) {
                    const double am = (acc_prev[dim * i + k] + acc[dim * i + k]) * 0.5; /* :311 */
                    vel[dim * i + k] += am * h;                                         /* :313 */
                }
                cap_speed(vel + dim * i, cfg->max_speed, dim);
            }
        }
    }
}

void orc_integrate(int64_t n, double* pos, double* vel, double* acc, double* acc_prev, const float* mass,
                   const uint8_t* pinned, const orc_config* cfg, float dt) {
    integrate_dim(n, pos, vel, acc, acc_prev, mass, pinned, cfg, dt, 2);
}

void orc_integrate3(int64_t n, double* pos, double* vel, double* acc, double* acc_prev, const float* mass,
                    const uint8_t* pinned, const orc_config* cfg, float dt) {
    integrate_dim(n, pos, vel, acc, acc_prev, mass, pinned, cfg, dt, 3);
}

/* One frame of the gravity + integrate systems (Physics.hpp:45-57), n_frames times. */
void orc_physics_frames(int64_t n, double* pos, double* vel, double* acc, double* acc_prev,
                        const float* mass, const uint8_t* pinned, const orc_config* cfg, float dt_eff,
                        int64_t n_frames) {
    for (int64_t f = 0; f < n_frames; ++f) {
        gravity_switch(n, pos, vel, mass, pinned, cfg, acc, 2);
        integrate_dim(n, pos, vel, acc, acc_prev, mass, pinned, cfg, dt_eff, 2);
    }
}

void orc_physics_frames3(int64_t n, double* pos, double* vel, double* acc, double* acc_prev,
                         const float* mass, const uint8_t* pinned, const orc_config* cfg, float dt_eff,
                         int64_t n_frames) {
    for (int64_t f = 0; f < n_frames; ++f) {
        gravity_switch(n, pos, vel, mass, pinned, cfg, acc, 3);
        integrate_dim(n, pos, vel, acc, acc_prev, mass, pinned, cfg, dt_eff, 3);
    }
}

/* compute_diagnostics(), Physics.hpp:117-177. Over ALL bodies, unfiltered (:120-121).
 * out[8] = kinetic, potential, energy, Px, Py, COMx, COMy, totalMass. Returns ok (1/0);
 * on a non-finite running sum the reference returns early with a zeroed struct (:123,:140-145). */
int orc_compute_diagnostics(int64_t n, const double* pos, const double* vel, const float* mass, double g,
                            double eps2, double* out) {
    memset(out, 0, 8 * sizeof(double));
    if (n == 0) return 1;
    double ke = 0.0, mt = 0.0, px = 0.0, py = 0.0, cx = 0.0, cy = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        const double m = (double)mass[i];
        const double vx = vel[2 * i], vy = vel[2 * i + 1];
        ke += 0.5 * m * (vx * vx + vy * vy); /* :134 */
        px += m * vx;
        py += m * vy;
        cx += m * pos[2 * i];
        cy += m * pos[2 * i + 1];
        mt += m;
        if (!(isfinite(ke) && isfinite(px) && isfinite(py) && isfinite(cx) && isfinite(cy) && isfinite(mt)))
            return 0;
    }
    double pe = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        for (int64_t j = i + 1; j < n; ++j) {
            const double dx = pos[2 * j] - pos[2 * i];
            const double dy = pos[2 * j + 1] - pos[2 * i + 1];
            const double r = sqrt(dx * dx + dy * dy + eps2);
            pe += -g * (double)mass[i] * (double)mass[j] / r; /* :154 */
            if (!isfinite(pe)) return 0;
        }
    }
    out[0] = ke;
    out[1] = pe;
    out[2] = ke + pe;
    out[3] = px;
    out[4] = py;
    out[5] = mt > 0.0 ? cx / mt : 0.0; /* :168 */
    out[6] = mt > 0.0 ? cy / mt : 0.0;
    out[7] = mt;
    for (int k = 0; k < 8; ++k)
        if (!isfinite(out[k])) return 0; /* :170-175 */
    return 1;
}

/* 3-D extension of the same sums: out[10] = KE, PE, E, Px, Py, Pz, COMx, COMy, COMz, M. */
int orc_compute_diagnostics3(int64_t n, const double* pos, const double* vel, const float* mass, double g,
                             double eps2, double* out) {
    memset(out, 0, 10 * sizeof(double));
    if (n == 0) return 1;
    double ke = 0.0, mt = 0.0, p[3] = {0, 0, 0}, c[3] = {0, 0, 0};
    for (int64_t i = 0; i < n; ++i) {
        const double m = (double)mass[i];
        const double* v = vel + 3 * i;
        ke += 0.5 * m * (v[0] * v[0] + v[1] * v[1] + v[2] * v[2]);
        for (int k = 0; k < 3; ++k) {
            p[k] += m * v[k];
            c[k] += m * pos[3 * i + k];
        }
        mt += m;
    }
    double pe = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        for (int64_t j = i + 1; j < n; ++j) {
            const double dx = pos[3 * j] - pos[3 * i];
            const double dy = pos[3 * j + 1] - pos[3 * i + 1];
            const double dz = pos[3 * j + 2] - pos[3 * i + 2];
            const double r = sqrt(dx * dx + dy * dy + dz * dz + eps2);
            pe += -g * (double)mass[i] * (double)mass[j] / r;
        }
    }
    out[0] = ke;
    out[1] = pe;
    out[2] = ke + pe;
    for (int k = 0; k < 3; ++k) {
        out[3 + k] = p[k];
        out[6 + k] = mt > 0.0 ? c[k] / mt : 0.0;
    }
    out[9] = mt;
    for (int k = 0; k < 10; ++k)
        if (!isfinite(out[k])) return 0;
    return 1;
}

/* zero_net_momentum(), Physics.hpp:66-85. */
void orc_zero_net_momentum(int64_t n, double* vel, const float* mass, const uint8_t* pinned) {
    double px = 0.0, py = 0.0, mt = 0.0;
    for (int64_t i = 0; i < n; ++i) {
        if (pinned && pinned[i]) {
            vel[2 * i] = 0.0;
            vel[2 * i + 1] = 0.0;
            continue;
        }
        px += (double)mass[i] * vel[2 * i];
        py += (double)mass[i] * vel[2 * i + 1];
        mt += (double)mass[i];
    }
    if (mt <= 0.0) return;
    const double v0x = px / mt, v0y = py / mt;
    for (int64_t i = 0; i < n; ++i) {
        if (pinned && pinned[i]) continue;
        vel[2 * i] -= v0x;
        vel[2 * i + 1] -= v0y;
    }
}

/* reset_scenario(), Physics.hpp:87-115: the built-in three-body seed. Every value goes
 * through float (raylib::Vector2) before widening (Math.hpp:12); constants from
 * src/core/Constants.hpp:47-51. Arrays must hold 3 bodies. */
void orc_seed_three_body(double g, double* pos, double* vel, float* mass, uint8_t* pinned) {
    const double central = 5.972e24, small = 7.342e22, radius = 3.844e8, cx = 0.0, cy = 0.0;
    const float v = (float)sqrt(g * central / radius); /* :109 */
    const float px[3] = {(float)cx, (float)(cx + radius), (float)(cx - radius)};
    const float vy[3] = {0.0f, v, -v};
    const float ms[3] = {(float)central, (float)small, (float)small};
    for (int i = 0; i < 3; ++i) {
        pos[2 * i] = (double)px[i];
        pos[2 * i + 1] = (double)(float)cy;
        vel[2 * i] = 0.0;
        vel[2 * i + 1] = (double)vy[i];
        mass[i] = ms[i];
        pinned[i] = 0;
    }
    orc_zero_net_momentum(3, vel, mass, pinned); /* :114 */
}

/* Detection phase of Collision::resolve, src/systems/Collision.hpp:63-84: bodies with finite
 * position and finite mass > 0 (:63-65); for i < j: dx = xj - xi, dy = yj - yi, rsum = ri + rj,
 * dist2 = dx*dx + dy*dy, overlap unless dist2 > rsum*rsum (:76-84). Returns the number of
 * overlapping pairs and the first one in visiting order ((-1,-1) if none). No merge is applied:
 * zero overlaps is exactly the case in which resolve() leaves the world untouched. */
int64_t orc_count_overlaps(int64_t n, const double* pos, const float* mass, const double* radius,
                           int64_t* first_pair) {
    int64_t count = 0;
    first_pair[0] = first_pair[1] = -1;
    for (int64_t i = 0; i < n; ++i) {
        if (!(isfinite(pos[2 * i]) && isfinite(pos[2 * i + 1]) && isfinite((double)mass[i])) || mass[i] <= 0.0f) continue;
        for (int64_t j = i + 1; j < n; ++j) {
            if (!(isfinite(pos[2 * j]) && isfinite(pos[2 * j + 1]) && isfinite((double)mass[j])) || mass[j] <= 0.0f) continue;
            const double dx = pos[2 * j] - pos[2 * i];
            const double dy = pos[2 * j + 1] - pos[2 * i + 1];
            const double rsum = radius[i] + radius[j];
            const double dist2 = dx * dx + dy * dy;
            if (dist2 > rsum * rsum) continue;
            if (count == 0) { first_pair[0] = i; first_pair[1] = j; }
            ++count;
        }
    }
    return count;
}

/* Collision::resolve in its default (merge) mode, src/systems/Collision.hpp:49-132, on flat arrays.
 * radius[i] is what Collision::radius_of gives for body i on entry (:28-35: the Radius component, else
 * the sphere of density[i]); density[i] is the Density component or Constants.hpp:34's 5510.
 * In/out: pos, vel, mass, pinned, radius (the survivor's new values, :115-121,128-132). Out: alive[i] = 0
 * for destructed bodies (:124-125), changed[i] = 1 for survivors whose components were rewritten (their
 * Acceleration and PrevAcceleration become 0, :118-119). Returns the number of merges.
 * Reproduced as they are: a row whose body dies keeps going with that body's snapshot — alive[i] is only
 * looked at when the row starts (:76) — so later partners can still merge INTO the dead body (and vanish
 * with it) or absorb its mass a second time. */
int64_t orc_resolve_collisions(int64_t n, double* pos, double* vel, float* mass, uint8_t* pinned, double* radius,
                               const double* density, uint8_t* alive, uint8_t* changed, int64_t* merge_log,
                               int64_t merge_log_cap) {
    const double pi = 3.141592653589793238462643383279502884; /* std::numbers::pi */
    int64_t* idx = (int64_t*)malloc(sizeof(int64_t) * (size_t)(n > 0 ? n : 1));
    int64_t m = 0, merges = 0;
    for (int64_t i = 0; i < n; ++i) {
        alive[i] = 1;
        changed[i] = 0;
        if (!(isfinite(pos[2 * i]) && isfinite(pos[2 * i + 1]) && isfinite((double)mass[i]))) continue; /* :63-64 */
        if (mass[i] <= 0.0f) continue;                                                                   /* :65 */
        idx[m++] = i;
    }
    if (m >= 2) { /* :70 */
        for (int64_t ii = 0; ii < m; ++ii) {
            const int64_t a = idx[ii];
            if (!alive[a]) continue; /* :76 */
            for (int64_t jj = ii + 1; jj < m; ++jj) {
                const int64_t b = idx[jj];
                if (!alive[b]) continue; /* :78 */
                const double dx = pos[2 * b] - pos[2 * a];
                const double dy = pos[2 * b + 1] - pos[2 * a + 1];
                const double rsum = radius[a] + radius[b];
                const double dist2 = dx * dx + dy * dy;
                if (dist2 > rsum * rsum) continue; /* :85 */
                const int a_survives = (mass[a] >= mass[b]) || pinned[b]; /* :94 */
                const int64_t s = a_survives ? a : b, d = a_survives ? b : a;
                if (pinned[d] && pinned[s]) continue; /* :97-104 */
                const double ms = (double)mass[s], md = (double)mass[d];
                const double M = ms + md;                                   /* :107 */
                const double px = ms * vel[2 * s] + md * vel[2 * d];        /* :108-109 */
                const double py = ms * vel[2 * s + 1] + md * vel[2 * d + 1];
                const double xx = (ms * pos[2 * s] + md * pos[2 * d]) / M;  /* :111-112 */
                const double xy = (ms * pos[2 * s + 1] + md * pos[2 * d + 1]) / M;
                mass[s] = (float)M;                                          /* :115,128 */
                vel[2 * s] = px / M;                                         /* :110,116 */
                vel[2 * s + 1] = py / M;
                pos[2 * s] = xx;
                pos[2 * s + 1] = xy;
                pinned[s] = (pinned[s] || pinned[d]) ? 1 : 0;                /* :120,131 */
                const double safe = (double)mass[s] > 1.0 ? (double)mass[s] : 1.0;
                radius[s] = cbrt((3.0 * safe) / (4.0 * pi * density[s]));    /* :37-46,121,132 */
                changed[s] = 1;
                alive[d] = 0;                                                /* :124-125 */
                if (merge_log && merges < merge_log_cap) { /* {survivor, absorbed} in the order of the merges */
                    merge_log[2 * merges] = s;
                    merge_log[2 * merges + 1] = d;
                }
                ++merges;
            }
        }
    }
    free(idx);
    return merges;
}

/* Collision::radius_of without a Radius component, Collision.hpp:30-34 (libm cbrt, like the reference). */
void orc_radius_from_mass(int64_t n, const float* mass, const double* density, double* radius) {
    const double pi = 3.141592653589793238462643383279502884;
    for (int64_t i = 0; i < n; ++i) {
        const double safe = (double)mass[i] > 1.0 ? (double)mass[i] : 1.0;
        radius[i] = cbrt((3.0 * safe) / (4.0 * pi * density[i]));
    }
}

/* libm's cbrt on an array: the reference's update_radius_from_mass (Collision.hpp:37-46) calls std::cbrt, so
 * the device restatement of it (csrc/merge.cuh::cbrt_glibc) is pinned against this machine's libm. */
void orc_cbrt(int64_t n, const double* in, double* out) {
    for (int64_t i = 0; i < n; ++i) out[i] = cbrt(in[i]);
}
