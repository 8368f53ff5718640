// PhysicsB200.hpp — the reference-side binding: what a raylib-nbody maintainer drops next to
// src/systems/Physics.hpp to run the physics step on a B200 through libnbody_b200.so.
//
// It is written against the reference's own types (Position/Velocity/Acceleration/
// PrevAcceleration/Mass/Pinned of src/components/Components.hpp:10-38, Config of
// src/core/Config.hpp:11-50, Physics::Diagnostics of src/systems/Physics.hpp:22-30) and replaces
//   Physics::register_systems     src/systems/Physics.hpp:32-64
//   Physics::compute_gravity      :183-261        Physics::integrate          :263-327
//   Physics::compute_diagnostics  :117-177        Physics::zero_net_momentum  :66-85
// and Collision::resolve (src/systems/Collision.hpp:49-132) by gather -> C-ABI call -> scatter. No reference
// code is copied: the reference keeps trails, UI and rendering.
//
// Device residency: the shim keeps a host mirror of what it last scattered.
// When a frame's gather equals the mirror nothing was edited from outside (UI, drag, merge,
// scenario load), the device state is still authoritative and no upload happens; any
// difference re-uploads everything. Row order is flecs iteration order of one w.each.
//
// Build: add include/ of this repo to the include path, link libnbody_b200.so.
#pragma once

#include <flecs.h>

#include <algorithm>
#include <cstdint>
#include <cstring>
#include <stdexcept>
#include <string>
#include <vector>

#include "components/Components.hpp"
#include "core/Config.hpp"
#include "core/Scenario.hpp"
#include "systems/Collision.hpp"
#include "systems/Physics.hpp"

#include "nbody_b200.h"

namespace nbody {

class PhysicsB200 {
public:
    using Diagnostics = Physics::Diagnostics;

    // Flat copy of the physics components of every body, in flecs iteration order.
    struct Bodies {
        std::vector<double> pos, vel, acc, acc_prev;  // 2 doubles per body (DVec2)
        std::vector<float> mass;
        std::vector<uint8_t> pinned;
        size_t size() const { return mass.size(); }
        bool same_inputs(const Bodies& o) const {
            // bitwise: NaN payloads and -0.0 count as edits too
            return mass.size() == o.mass.size() && bytes_equal(pos, o.pos) && bytes_equal(vel, o.vel) &&
                   bytes_equal(mass, o.mass) && bytes_equal(pinned, o.pinned);
        }
        template <class T>
        static bool bytes_equal(const std::vector<T>& a, const std::vector<T>& b) {
            return a.size() == b.size() && (a.empty() || std::memcmp(a.data(), b.data(), a.size() * sizeof(T)) == 0);
        }
    };

    // One per world; stored as a flecs singleton by register_systems.
    struct Device {
        nb_ctx* ctx = nullptr;
        Bodies mirror;            // what the device holds, as last scattered / uploaded
        bool resident = false;    // mirror matches device state
        nb_config applied{};
        bool config_applied = false;
        // NB_SOLVER_DIRECT: always the all-pairs sum (Config::bh_threshold ignored). NB_SOLVER_REFERENCE:
        // honour Config::bh_threshold / bh_theta exactly like Physics.hpp:212 — above the threshold the
        // reference's float Barnes-Hut quadtree, bit-identical, so frames equal the reference's at its
        // DEFAULT settings (bh_threshold = 100).
        int solver = NB_SOLVER_DIRECT;
        int applied_solver = -1, applied_bh_threshold = 0;
        float applied_bh_theta = 0.0f;
        int n_devices = 1;
        long long uploads = 0, frames = 0;
        long long collision_checks = 0, collision_frames = 0, merges = 0;  // device passes / frames with a merge / merges
        // Verlet: the kick pass of frame k also accumulates the potential at the positions frame k+1 starts
        // from, so its fused diagnostics ARE frame k+1's frame-start diagnostics unless something was edited
        bool diag_cached = false;
        nb_diag cached_diag{};
        long long diag_reused = 0;
    };

    static void check(int status, const char* what) {
        if (status != NB_OK) throw std::runtime_error(std::string(what) + ": " + nb_last_error());
    }

    static Bodies gather(const flecs::world& w) {
        Bodies b;
        w.each([&](const Position& p, const Velocity& v, const Acceleration& a, const PrevAcceleration& a0,
                   const Mass& m, const Pinned& pin) {
            b.pos.push_back(p.value.x); b.pos.push_back(p.value.y);
            b.vel.push_back(v.value.x); b.vel.push_back(v.value.y);
            b.acc.push_back(a.value.x); b.acc.push_back(a.value.y);
            b.acc_prev.push_back(a0.value.x); b.acc_prev.push_back(a0.value.y);
            b.mass.push_back(m.value);
            b.pinned.push_back(pin.value ? 1 : 0);
        });
        return b;
    }

    // Write the device state back into the components (Physics.hpp:260 and the in-place writes
    // of :276-323). `what`: bit 0 pos+vel, bit 1 acc, bit 2 acc_prev.
    static void scatter(const flecs::world& w, const Bodies& b, int what) {
        size_t i = 0;
        w.each([&](Position& p, Velocity& v, Acceleration& a, PrevAcceleration& a0, const Mass&, const Pinned&) {
            if (what & 1) {
                p.value = DVec2{b.pos[2 * i], b.pos[2 * i + 1]};
                v.value = DVec2{b.vel[2 * i], b.vel[2 * i + 1]};
            }
            if (what & 2) a.value = DVec2{b.acc[2 * i], b.acc[2 * i + 1]};
            if (what & 4) a0.value = DVec2{b.acc_prev[2 * i], b.acc_prev[2 * i + 1]};
            ++i;
        });
    }

    static void apply_config(Device& d, const Config& cfg) {
        nb_config c;
        std::memset(&c, 0, sizeof(c));  // padding too: the struct is compared bytewise below
        c.g = cfg.g;
        c.softening = cfg.softening;
        c.max_speed = cfg.max_speed;
        c.integrator = cfg.integrator == 0 ? NB_INTEGRATOR_EULER : NB_INTEGRATOR_VERLET;  // Physics.hpp:273,294
        c.max_substep = cfg.max_substep;
        c.max_substeps_per_frame = cfg.max_substeps_per_frame;
        if (d.applied_solver != d.solver || d.applied_bh_threshold != cfg.bh_threshold ||
            std::memcmp(&d.applied_bh_theta, &cfg.bh_theta, sizeof(float)) != 0) {
            check(nb_set_solver(d.ctx, d.solver, cfg.bh_threshold, cfg.bh_theta), "nb_set_solver");
            d.applied_solver = d.solver;
            d.applied_bh_threshold = cfg.bh_threshold;
            d.applied_bh_theta = cfg.bh_theta;
        }
        if (d.config_applied && std::memcmp(&c, &d.applied, sizeof(c)) == 0) return;
        check(nb_set_config(d.ctx, &c), "nb_set_config");
        d.diag_cached = false;
        d.applied = c;
        d.config_applied = true;
    }

    // Make the device hold the world's current bodies; returns the gathered copy.
    static Bodies sync(const flecs::world& w, Device& d) {
        Bodies now = gather(w);
        if (!(d.resident && now.same_inputs(d.mirror))) {
            // external edit (or first frame): Acceleration is not trusted to equal a(x), exactly
            // like the reference's gravity system recomputing it every frame (Physics.hpp:45-48)
            check(nb_upload_state(d.ctx, (int64_t)now.size(), 2, now.pos.data(), now.vel.data(), nullptr,
                                  now.mass.data(), now.pinned.data()),
                  "nb_upload_state");
            d.uploads++;
            d.mirror = now;
            d.resident = true;
            d.diag_cached = false;
        }
        return now;
    }

    static void pull(Device& d, Bodies& b, bool with_prev) {
        check(nb_download_state(d.ctx, 2, b.pos.data(), b.vel.data(), b.acc.data(),
                                with_prev ? b.acc_prev.data() : nullptr),
              "nb_download_state");
    }

    // ---- the reference statics, one by one (same names, arguments and side effects) --------

    // Physics::compute_diagnostics, Physics.hpp:117-177. G and eps2 come from Config on the
    // device side (the reference's callers always pass cfg->g and cfg->softening^2, :38-40).
    static bool compute_diagnostics(const flecs::world& w, Device& d, Diagnostics& out) {
        out = Diagnostics{};
        auto* cfg = w.get_mut<Config>();
        apply_config(d, *cfg);
        sync(w, d);
        nb_diag r;
        if (d.diag_cached) {
            r = d.cached_diag;
            d.diag_reused++;
        } else {
            check(nb_compute_diagnostics(d.ctx, &r), "nb_compute_diagnostics");
        }
        if (!r.ok) {  // Physics.hpp:140-145,155-159,170-175: auto-pause, zeroed struct
            cfg->paused = true;
            out.ok = false;
            return false;
        }
        out.kinetic = r.kinetic;
        out.potential = r.potential;
        out.energy = r.energy;
        out.momentum = DVec2{r.momentum[0], r.momentum[1]};
        out.com = DVec2{r.com[0], r.com[1]};
        out.totalMass = r.total_mass;
        out.ok = true;
        return true;
    }

    // Physics::compute_gravity, Physics.hpp:183-261 (the direct sum, or the reference's switch when
    // Device::solver == NB_SOLVER_REFERENCE).
    static void compute_gravity(const flecs::world& w, Device& d) {
        apply_config(d, *w.get<Config>());
        Bodies b = sync(w, d);
        if (b.size() == 0) return;
        check(nb_compute_gravity(d.ctx), "nb_compute_gravity");
        pull(d, b, false);
        scatter(w, b, 2);
        d.mirror.acc = b.acc;
    }

    // Physics::integrate, Physics.hpp:263-327: advances with whatever Acceleration holds.
    static void integrate(const flecs::world& w, Device& d, float dt) {
        const Config& cfg = *w.get<Config>();
        apply_config(d, cfg);
        Bodies b = gather(w);
        if (b.size() == 0) return;
        check(nb_upload_state(d.ctx, (int64_t)b.size(), 2, b.pos.data(), b.vel.data(), b.acc.data(), b.mass.data(),
                              b.pinned.data()),
              "nb_upload_state");
        d.uploads++;
        d.diag_cached = false;
        check(nb_integrate(d.ctx, dt), "nb_integrate");
        const bool verlet = cfg.integrator != 0;
        pull(d, b, verlet);
        scatter(w, b, verlet ? 7 : 1);
        d.mirror = b;
        d.resident = true;
    }

    // Physics::zero_net_momentum, Physics.hpp:66-85.
    static void zero_net_momentum(const flecs::world& w, Device& d) {
        Bodies b = sync(w, d);
        if (b.size() == 0) return;
        d.diag_cached = false;
        check(nb_zero_net_momentum(d.ctx), "nb_zero_net_momentum");
        pull(d, b, false);
        scatter(w, b, 1);
        d.mirror.vel = b.vel;
    }

    // Collision::resolve (Collision.hpp:49-132, merge mode) through nb_resolve_collisions: every pair test runs
    // on the device, the merges are replayed in the reference's order (bit-identical outcome, including the
    // rows that keep running after their body died), and the outcome is applied to the components with the
    // same calls the reference makes: immediate writes through get_mut, deferred destruct (:115-124) — so the
    // diagnostics that follow in the same system see what they see in the reference (survivors updated, the
    // destructed bodies still present until the system ends).
    static void resolve_collisions(const flecs::world& w, Device& d) {
        apply_config(d, *w.get<Config>());
        Bodies b = sync(w, d);
        const size_t n = b.size();
        if (n < 2) return;  // Collision.hpp:70
        if (d.n_devices > 1) {
            // the collision pass is a single-GPU operation: a multi-GPU world must not need it
            bool any = false;
            w.each([&](const flecs::entity e, const Position&, const Mass& m) {
                any = any || nbody::systems::Collision::radius_of(e, m) > 0.0;
            });
            if (any) throw std::runtime_error("PhysicsB200: a multi-GPU context has no collision pass; give bodies Radius{0}");
            return;
        }
        std::vector<flecs::entity> ents;
        std::vector<double> radius, density;
        ents.reserve(n); radius.reserve(n); density.reserve(n);
        w.each([&](const flecs::entity e, const Position&, const Velocity&, const Acceleration&,
                   const PrevAcceleration&, const Mass& m, const Pinned&) {
            ents.push_back(e);
            radius.push_back(nbody::systems::Collision::radius_of(e, m));  // Collision.hpp:28-35
            double rho = nbody::constants::body_density;                   // :31-32
            if (const auto* dd = e.get<Density>()) rho = dd->value;
            density.push_back(rho);
        });
        std::vector<uint8_t> alive(n), changed(n), pinned(n);
        std::vector<double> pos(2 * n), vel(2 * n), rad(n);
        std::vector<float> mass(n);
        int64_t merges = 0;
        std::vector<int64_t> log(2 * 4 * n);
        for (;;) {
            check(nb_resolve_collisions(d.ctx, radius.data(), density.data(), &merges, alive.data(), changed.data(),
                                        pos.data(), vel.data(), mass.data(), pinned.data(), rad.data(), log.data(),
                                        (int64_t)(log.size() / 2)),
                  "nb_resolve_collisions");
            if ((size_t)merges * 2 <= log.size()) break;
            log.resize(2 * (size_t)merges);  // more merges than bodies x 4 (rows of dead bodies): once more with room
        }
        d.collision_checks++;
        if (merges == 0) return;
        d.collision_frames++;
        d.merges += merges;
        // Per merge, in the reference's order, the very calls it makes (:115-124). The component writes go through
        // get_mut and take effect at once, so writing the survivor's FINAL values at each of its merges leaves the
        // same components; but every get_mut / set / destruct inside a system also enters flecs' command queue,
        // and the queue's order decides the order of the table moves and swap-removes at the end of the system —
        // i.e. the row order of the bodies that are left. Same calls in the same order => same rows.
        for (int64_t k = 0; k < merges; ++k) {
            const size_t i = (size_t)log[2 * k];
            flecs::entity sv = ents[i], dead = ents[(size_t)log[2 * k + 1]];
            if (auto* mp = sv.get_mut<Mass>()) mp->value = mass[i];
            if (auto* vp = sv.get_mut<Velocity>()) vp->value = DVec2{vel[2 * i], vel[2 * i + 1]};
            if (auto* pp = sv.get_mut<Position>()) pp->value = DVec2{pos[2 * i], pos[2 * i + 1]};
            if (auto* ap = sv.get_mut<Acceleration>()) ap->value = DVec2{0.0, 0.0};
            if (auto* a0 = sv.get_mut<PrevAcceleration>()) a0->value = DVec2{0.0, 0.0};
            if (auto* pinp = sv.get_mut<Pinned>()) pinp->value = pinned[i] != 0;
            if (auto* rp = sv.get_mut<Radius>()) rp->value = rad[i]; else sv.set<Radius>({rad[i]});  // :42-46
            dead.destruct();
        }
    }

    // ---- the frame, Physics.hpp:32-64 -------------------------------------------------------
    // sysA = reference system 1 (collision pass and diagnostics on the device),
    // sysB = reference systems 2+3 fused: one nb_step evaluates gravity and integrates, which is
    // what the gravity system followed by integrate() computes. Trails (system 4) stay with the
    // reference's renderer.
    static void register_systems(const flecs::world& w, int device = 0, int solver = NB_SOLVER_DIRECT) {
        register_systems(w, std::vector<int>{device}, solver);
    }

    // Several GPUs behind the same single-threaded flecs systems (the reference is one process, main.cpp:100-143):
    // nb_create_multi partitions the targets over `devices` and exchanges packed positions through peer memory.
    // Multi-GPU contexts run the all-pairs path only: no collision pass (bodies must carry Radius{0}) and the
    // direct solver.
    static void register_systems(const flecs::world& w, const std::vector<int>& devices, int solver = NB_SOLVER_DIRECT) {
        Device d;
        d.solver = solver;
        check(nb_create_multi(devices.data(), (int)devices.size(), &d.ctx), "nb_create_multi");
        d.n_devices = nb_num_devices(d.ctx);
        w.set<Device>(d);

        w.system<>().kind(flecs::OnUpdate).iter([&](flecs::iter&) {
            auto* cfg = w.get<Config>();
            if (!cfg || cfg->paused) return;
            resolve_collisions(w, *w.get_mut<Device>());
            Diagnostics diag{};
            diag.ok = compute_diagnostics(w, *w.get_mut<Device>(), diag);
            w.set<Diagnostics>(diag);
        });

        w.system<>().kind(flecs::OnUpdate).iter([&](const flecs::iter& it) {
            const Config& cfg = *w.get<Config>();
            if (cfg.paused) return;
            Device& dev = *w.get_mut<Device>();
            apply_config(dev, cfg);
            const float baseDt = cfg.use_fixed_dt ? cfg.fixed_dt : (float)it.delta_time();
            const float dtEff = baseDt * std::max(0.0f, cfg.time_scale);  // Physics.hpp:54-55
            Bodies b = sync(w, dev);  // merges flushed at the end of sysA show up here as an edit
            if (b.size() == 0) return;
            const bool verlet = cfg.integrator != 0;
            dev.diag_cached = false;
            if (verlet) {
                check(nb_step_with_diagnostics(dev.ctx, dtEff, 1), "nb_step_with_diagnostics");
                check(nb_step_diagnostics(dev.ctx, &dev.cached_diag), "nb_step_diagnostics");
                dev.diag_cached = true;
            } else {
                check(nb_step(dev.ctx, dtEff, 1), "nb_step");
            }
            pull(dev, b, verlet);
            scatter(w, b, verlet ? 7 : 3);
            dev.mirror = b;
            dev.frames++;
        });
    }

    // ---- scenarios on disk (the reference keeps them in memory only, src/core/Scenario.hpp:14-128) -----------
    // A Scenario's bodies (pos, vel, mass, pinned) and the physics part of its config subset, in the C-ABI's
    // ".nbs" format (nb_scenario_write), which scenarios.py reads and writes too. Render-only fields of a
    // Scenario (name, tags, tint, meter_to_pixel, draw_*, trail_max, radius_scale) are UI state and stay out.
    static void save_scenario(const Scenario& sc, const std::string& path) {
        const size_t n = sc.bodies.size();
        std::vector<double> pos(2 * n), vel(2 * n);
        std::vector<float> mass(n);
        std::vector<uint8_t> pinned(n);
        for (size_t i = 0; i < n; ++i) {
            pos[2 * i] = sc.bodies[i].pos.x; pos[2 * i + 1] = sc.bodies[i].pos.y;
            vel[2 * i] = sc.bodies[i].vel.x; vel[2 * i + 1] = sc.bodies[i].vel.y;
            mass[i] = sc.bodies[i].mass;
            pinned[i] = sc.bodies[i].pinned ? 1 : 0;
        }
        nb_scenario_config c{};
        c.physics.g = sc.g; c.physics.softening = sc.softening; c.physics.max_speed = sc.max_speed;
        c.physics.integrator = sc.integrator; c.physics.max_substep = sc.max_substep;
        c.physics.max_substeps_per_frame = sc.max_substeps_per_frame;
        c.bh_threshold = sc.bh_threshold; c.bh_theta = sc.bh_theta;
        c.use_fixed_dt = sc.use_fixed_dt ? 1 : 0; c.fixed_dt = sc.fixed_dt; c.time_scale = sc.time_scale;
        check(nb_scenario_write(path.c_str(), (int64_t)n, 2, pos.data(), vel.data(), nullptr, nullptr, mass.data(),
                                pinned.data(), &c, nullptr, 0, 0),
              "nb_scenario_write");
    }

    // snapshot_from_world (Scenario.hpp:53-82) straight to a file.
    static void save_scenario(const flecs::world& w, const std::string& path) {
        save_scenario(snapshot_from_world(w, "", ""), path);
    }

    // A Scenario from a file; apply it with the reference's own apply_scenario_to_world (Scenario.hpp:107-128):
    // the binding notices the new bodies at the next frame (mirror mismatch) and uploads them.
    static Scenario load_scenario(const std::string& path) {
        int64_t n = 0;
        int dim = 0;
        nb_scenario_config c{};
        check(nb_scenario_read_header(path.c_str(), &n, &dim, nullptr, &c, nullptr, nullptr), "nb_scenario_read_header");
        if (dim != 2) throw std::runtime_error("PhysicsB200::load_scenario: the reference's worlds are 2-D");
        std::vector<double> pos(2 * (size_t)n), vel(2 * (size_t)n);
        std::vector<float> mass((size_t)n);
        std::vector<uint8_t> pinned((size_t)n);
        check(nb_scenario_read(path.c_str(), pos.data(), vel.data(), nullptr, nullptr, mass.data(), pinned.data(), nullptr),
              "nb_scenario_read");
        Scenario sc{};
        sc.bodies.resize((size_t)n);
        for (size_t i = 0; i < (size_t)n; ++i) {
            sc.bodies[i].pos = DVec2{pos[2 * i], pos[2 * i + 1]};
            sc.bodies[i].vel = DVec2{vel[2 * i], vel[2 * i + 1]};
            sc.bodies[i].mass = mass[i];
            sc.bodies[i].pinned = pinned[i] != 0;
        }
        sc.g = c.physics.g; sc.softening = c.physics.softening; sc.max_speed = c.physics.max_speed;
        sc.integrator = c.physics.integrator; sc.max_substep = c.physics.max_substep;
        sc.max_substeps_per_frame = c.physics.max_substeps_per_frame;
        sc.bh_threshold = c.bh_threshold; sc.bh_theta = c.bh_theta;
        sc.use_fixed_dt = c.use_fixed_dt != 0; sc.fixed_dt = c.fixed_dt; sc.time_scale = c.time_scale;
        return sc;
    }

    static void shutdown(const flecs::world& w) {
        if (auto* d = w.get_mut<Device>(); d && d->ctx) {
            nb_destroy(d->ctx);
            d->ctx = nullptr;
        }
    }
};

}  // namespace nbody
