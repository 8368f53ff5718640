// oracle/ref_shim.cpp — TEST INFRASTRUCTURE, not product code.
//
// Thin C entry points around the UNMODIFIED reference physics, compiled from the
// sources where they lie under /root/reference (nothing is copied into this repo).
// Built only by oracle/Makefile into oracle/_ref/libnbref.so. Used by tests/, by
// __graft_entry__.smoke() and by bench.py's cpu_baseline / --impl reference legs
// as the checker and the CPU baseline — never on the product path.
//
// What it wraps (reference file:line):
//   Physics::compute_gravity      src/systems/Physics.hpp:183-261 (private; reached via
//                                 the `#define private public` below, no reference file is modified)
//   Physics::integrate            src/systems/Physics.hpp:263-327 (private)
//   Physics::compute_diagnostics  src/systems/Physics.hpp:117-177
//   Physics::zero_net_momentum    src/systems/Physics.hpp:66-85
//   Physics::reset_scenario       src/systems/Physics.hpp:87-115
//   Physics::register_systems     src/systems/Physics.hpp:32-64  (full world.progress() frame)
//
// Mandatory oracle settings: bh_threshold >= N so the direct-sum
// branch (Physics.hpp:231-258) runs, Radius{0} on every body so Collision::resolve
// never merges, use_fixed_dt=true, draw_trails=false.

#include <algorithm>
#include <cmath>
#include <climits>
#include <cstdint>
#include <cstring>
#include <numbers>
#include <tuple>
#include <vector>

#include <flecs.h>
#include <raylib-cpp.hpp>
#include <raymath.h>

#define private public
#include "systems/Physics.hpp"
#undef private

namespace {

struct RefWorld {
    flecs::world w;
    std::vector<flecs::entity> bodies;  // creation order == flecs table row order
    bool systems_registered = false;
};

void collect_bodies(RefWorld* rw) {
    rw->bodies.clear();
    rw->w.each([&](const flecs::entity e, Position&) { rw->bodies.push_back(e); });
}

}  // namespace

extern "C" {

void* ref_world_create(void) {
    auto* rw = new RefWorld();
    Config cfg{};
    cfg.bh_threshold = INT_MAX;  // force the direct-sum branch (Physics.hpp:212)
    cfg.use_fixed_dt = true;
    cfg.draw_trails = false;
    rw->w.set<Config>(cfg);
    return rw;
}

void ref_world_destroy(void* h) { delete static_cast<RefWorld*>(h); }

// Physics fields of Config (src/core/Config.hpp:13-29).
void ref_set_config(void* h, double g, float softening, float max_speed, int integrator, float fixed_dt,
                    float time_scale, float max_substep, int max_substeps_per_frame) {
    auto* rw = static_cast<RefWorld*>(h);
    Config* cfg = rw->w.get_mut<Config>();
    cfg->g = g;
    cfg->softening = softening;
    cfg->max_speed = max_speed;
    cfg->integrator = integrator;
    cfg->fixed_dt = fixed_dt;
    cfg->time_scale = time_scale;
    cfg->max_substep = max_substep;
    cfg->max_substeps_per_frame = max_substeps_per_frame;
    cfg->paused = false;
}

// The gravity solver switch (Config.hpp:17-18, Physics.hpp:212): Barnes-Hut above bh_threshold bodies.
// ref_world_create() sets INT_MAX (direct sum); the reference's own default is 100 / 0.5.
void ref_set_solver(void* h, int bh_threshold, float bh_theta) {
    Config* cfg = static_cast<RefWorld*>(h)->w.get_mut<Config>();
    cfg->bh_threshold = bh_threshold;
    cfg->bh_theta = bh_theta;
}

void ref_get_solver_defaults(int* bh_threshold, float* bh_theta) {
    Config cfg{};
    *bh_threshold = cfg.bh_threshold;
    *bh_theta = cfg.bh_theta;
}

void ref_get_defaults(double* g, float* softening, float* max_speed, int* integrator, float* fixed_dt,
                      float* time_scale, float* max_substep, int* max_substeps_per_frame) {
    Config cfg{};
    *g = cfg.g;
    *softening = cfg.softening;
    *max_speed = cfg.max_speed;
    *integrator = cfg.integrator;
    *fixed_dt = cfg.fixed_dt;
    *time_scale = cfg.time_scale;
    *max_substep = cfg.max_substep;
    *max_substeps_per_frame = cfg.max_substeps_per_frame;
}

int ref_paused(void* h) { return static_cast<RefWorld*>(h)->w.get<Config>()->paused ? 1 : 0; }

// Replace all bodies. pos/vel: interleaved x,y doubles (DVec2 layout, Math.hpp:6-9);
// mass: f32 (Components.hpp:22-24); pinned: u8 (Components.hpp:36-38).
void ref_load_bodies(void* h, int64_t n, const double* pos, const double* vel, const float* mass,
                     const uint8_t* pinned) {
    auto* rw = static_cast<RefWorld*>(h);
    for (auto& e : rw->bodies) e.destruct();
    rw->bodies.clear();
    rw->bodies.reserve(static_cast<size_t>(n));
    for (int64_t i = 0; i < n; ++i) {
        auto e = rw->w.entity()
                     .set<Position>({DVec2{pos[2 * i], pos[2 * i + 1]}})
                     .set<Velocity>({DVec2{vel[2 * i], vel[2 * i + 1]}})
                     .set<Acceleration>({DVec2{0.0, 0.0}})
                     .set<PrevAcceleration>({DVec2{0.0, 0.0}})
                     .set<Mass>({mass[i]})
                     .set<Pinned>({pinned ? pinned[i] != 0 : false})
                     .set<Radius>({0.0});
        rw->bodies.push_back(e);
    }
}

// Explicit per-body Radius components (Components.hpp:33-35), then the reference's collision pass
// (Collision.hpp:49-203, merge mode). Returns the number of bodies left; used to pin the
// oracle's detection: resolve() changes nothing exactly when no pair overlaps.
int64_t ref_collision_resolve(void* h, const double* radius) {
    auto* rw = static_cast<RefWorld*>(h);
    for (size_t i = 0; i < rw->bodies.size(); ++i) rw->bodies[i].set<Radius>({radius[i]});
    rw->w.defer_begin();  // as inside a flecs system: destruct() takes effect at defer_end (flecs.c:64517,64556)
    nbody::systems::Collision::resolve(rw->w);
    rw->w.defer_end();
    collect_bodies(rw);
    return static_cast<int64_t>(rw->bodies.size());
}

// The same pass with everything it changed reported per ORIGINAL body (creation order of ref_load_bodies):
// radius_in < 0 leaves the body without a Radius component (radius from mass and density, Collision.hpp:30-34),
// density_in <= 0 leaves it without a Density component (Constants.hpp:34 applies). Outputs: alive flag and the
// components of the bodies that are still alive; radius_out = Collision::radius_of after the pass.
// visit_order[k] = original index of the k-th body the pass visits: flecs iterates table by table, and bodies
// with and without a Radius (or Density) component live in different tables, so this is NOT creation order
// in general. The caller restates the pass on the arrays permuted this way.
int64_t ref_collision_resolve_full(void* h, const double* radius_in, const double* density_in, uint8_t* alive,
                                   double* pos, double* vel, double* acc, float* mass, uint8_t* pinned,
                                   double* radius_out, int64_t* visit_order) {
    auto* rw = static_cast<RefWorld*>(h);
    std::vector<flecs::entity> orig = rw->bodies;
    for (size_t i = 0; i < orig.size(); ++i) {
        if (radius_in[i] >= 0.0) orig[i].set<Radius>({radius_in[i]}); else orig[i].remove<Radius>();
        if (density_in && density_in[i] > 0.0) orig[i].set<Density>({density_in[i]});
    }
    {
        int64_t k = 0;
        rw->w.each([&](const flecs::entity e, const Position&, const Velocity&, const Mass&, const Pinned&) {
            for (size_t i = 0; i < orig.size(); ++i)
                if (orig[i] == e) { visit_order[k++] = (int64_t)i; break; }
        });
    }
    rw->w.defer_begin();
    nbody::systems::Collision::resolve(rw->w);
    rw->w.defer_end();
    int64_t left = 0;
    for (size_t i = 0; i < orig.size(); ++i) {
        alive[i] = orig[i].is_alive() ? 1 : 0;
        if (!alive[i]) continue;
        ++left;
        const auto* p = orig[i].get<Position>();
        const auto* v = orig[i].get<Velocity>();
        const auto* a = orig[i].get<Acceleration>();
        pos[2 * i] = p->value.x; pos[2 * i + 1] = p->value.y;
        vel[2 * i] = v->value.x; vel[2 * i + 1] = v->value.y;
        acc[2 * i] = a->value.x; acc[2 * i + 1] = a->value.y;
        mass[i] = orig[i].get<Mass>()->value;
        pinned[i] = orig[i].get<Pinned>()->value ? 1 : 0;
        radius_out[i] = nbody::systems::Collision::radius_of(orig[i], *orig[i].get<Mass>());
    }
    collect_bodies(rw);
    return left;
}

// Built-in three-body seed with momentum zeroing (Physics.hpp:87-115). Radius{0} is
// added afterwards so a later ref_progress() cannot merge bodies.
int64_t ref_reset_scenario(void* h) {
    auto* rw = static_cast<RefWorld*>(h);
    nbody::Physics::reset_scenario(rw->w);
    collect_bodies(rw);
    for (auto& e : rw->bodies) e.set<Radius>({0.0});
    collect_bodies(rw);
    return static_cast<int64_t>(rw->bodies.size());
}

int64_t ref_num_bodies(void* h) { return static_cast<int64_t>(static_cast<RefWorld*>(h)->bodies.size()); }

void ref_read_bodies(void* h, double* pos, double* vel, double* acc, float* mass, uint8_t* pinned) {
    auto* rw = static_cast<RefWorld*>(h);
    int64_t i = 0;
    for (auto& e : rw->bodies) {
        if (pos) {
            const auto* p = e.get<Position>();
            pos[2 * i] = p->value.x;
            pos[2 * i + 1] = p->value.y;
        }
        if (vel) {
            const auto* v = e.get<Velocity>();
            vel[2 * i] = v->value.x;
            vel[2 * i + 1] = v->value.y;
        }
        if (acc) {
            const auto* a = e.get<Acceleration>();
            acc[2 * i] = a->value.x;
            acc[2 * i + 1] = a->value.y;
        }
        if (mass) mass[i] = e.get<Mass>()->value;
        if (pinned) pinned[i] = e.get<Pinned>()->value ? 1 : 0;
        ++i;
    }
}

void ref_zero_net_momentum(void* h) { nbody::Physics::zero_net_momentum(static_cast<RefWorld*>(h)->w); }

void ref_compute_gravity(void* h) { nbody::Physics::compute_gravity(static_cast<RefWorld*>(h)->w); }

void ref_integrate(void* h, float dt) { nbody::Physics::integrate(static_cast<RefWorld*>(h)->w, dt); }

// dtEff exactly as the integrate system computes it (Physics.hpp:54-55) with use_fixed_dt.
float ref_dt_eff(void* h) {
    const Config& cfg = *static_cast<RefWorld*>(h)->w.get<Config>();
    return cfg.fixed_dt * std::max(0.0f, cfg.time_scale);
}

// One frame of the gravity + integrate systems (sys2 + sys3 of Physics.hpp:45-57)
// without Collision/diagnostics/trails: what "one physics step" means on the hot path.
void ref_physics_frame(void* h) {
    auto* rw = static_cast<RefWorld*>(h);
    const Config& cfg = *rw->w.get<Config>();
    if (cfg.paused) return;
    nbody::Physics::compute_gravity(rw->w);
    nbody::Physics::integrate(rw->w, ref_dt_eff(h));
}

// out[8] = kinetic, potential, energy, Px, Py, COMx, COMy, totalMass (Physics.hpp:22-30).
int ref_compute_diagnostics(void* h, double* out) {
    auto* rw = static_cast<RefWorld*>(h);
    const Config& cfg = *rw->w.get<Config>();
    nbody::Physics::Diagnostics d{};
    const double eps2 = static_cast<double>(cfg.softening) * static_cast<double>(cfg.softening);
    const bool ok = nbody::Physics::compute_diagnostics(rw->w, cfg.g, eps2, d);
    out[0] = d.kinetic;
    out[1] = d.potential;
    out[2] = d.energy;
    out[3] = d.momentum.x;
    out[4] = d.momentum.y;
    out[5] = d.com.x;
    out[6] = d.com.y;
    out[7] = d.totalMass;
    return ok ? 1 : 0;
}

// The reference's full frame through its own flecs pipeline (collision + diagnostics,
// gravity, integrate, trails): world.progress(fixed_dt), main.cpp:137.
void ref_progress(void* h) {
    auto* rw = static_cast<RefWorld*>(h);
    if (!rw->systems_registered) {
        nbody::Physics::register_systems(rw->w);
        rw->systems_registered = true;
    }
    rw->w.progress(rw->w.get<Config>()->fixed_dt);
}

int ref_read_frame_diagnostics(void* h, double* out) {
    auto* rw = static_cast<RefWorld*>(h);
    const auto* d = rw->w.get<nbody::Physics::Diagnostics>();
    if (!d) return -1;
    out[0] = d->kinetic;
    out[1] = d->potential;
    out[2] = d->energy;
    out[3] = d->momentum.x;
    out[4] = d->momentum.y;
    out[5] = d->com.x;
    out[6] = d->com.y;
    out[7] = d->totalMass;
    return d->ok ? 1 : 0;
}

}  // extern "C"
