#!/usr/bin/env python
"""bench.py — body-body interactions/s of the physics step on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU code, same metric

    python bench.py --config 5 --gpus N                      # 2^22-body two-galaxy collision (strong scaling, drift)

A step = one frame of config 4 — the 2^20-body Plummer sphere, Velocity Verlet — issued the way the flecs
binding issues it: nb_step(ctx, dt, 1). In steady state that is ONE kernel launch per GPU (two when
sharded: own sources, then remote sources): the force evaluation over N^2 body-body interactions with
kick, the next frame's drift (into shadow positions, committed by the next call) and its packed stream
in the epilogue. Targets are partitioned over the ranks (one process per GPU); the all-gather of the
packed stream is issued behind each step and overlaps the next step's own-sources launch. The line
carries a parity record against the unmodified reference's full-size accelerations at every rank count
and exits non-zero when it fails. One JSON line on stdout from rank 0; everything else goes to stderr.
"""
import argparse
import json
import os
import subprocess
import sys
import threading
import time

import numpy as np

ROOT = os.path.dirname(os.path.abspath(__file__))
sys.path.insert(0, ROOT)

FLOPS_PER_INTERACTION = 20.0
FP32_LANES_PER_SM = 128
METRIC = "body-body interactions/s"
UNIT = "interactions/s"


_JSON_OUT = None


def emit(line):
    out = _JSON_OUT or sys.stdout
    out.write(json.dumps(line) + "\n")
    out.flush()


def log(*a):
    print(*a, file=sys.stderr, flush=True)


def workload(n, dim=3):
    import __graft_entry__ as ge
    pkg = ge.load_package()
    pos, vel, mass = pkg.scenarios.plummer(n, seed=1234, dim=dim)
    eps = 3.844e8 / np.sqrt(n)
    dt = 1e-3 * pkg.scenarios.crossing_time(n)
    return pkg, pos, vel, mass, eps, dt


class ClockSampler:
    """nvidia-smi clocks and throttle reasons during the timed region (B200_PROFILING.md)."""
    Q = ("index,clocks.sm,clocks.max.sm,power.draw,clocks_event_reasons.active,"
         "clocks_event_reasons.hw_slowdown,clocks_event_reasons.hw_thermal_slowdown,"
         "clocks_event_reasons.sw_thermal_slowdown,clocks_event_reasons.sw_power_cap")

    def __init__(self, gpu_index=0):
        self.rows = []
        self.proc = None
        self.idx = gpu_index

    def start(self):
        try:
            self.proc = subprocess.Popen(["nvidia-smi", f"--query-gpu={self.Q}", "--format=csv,noheader,nounits",
                                          "-lms", "100", "-i", str(self.idx)], stdout=subprocess.PIPE, text=True)
            threading.Thread(target=self._pump, daemon=True).start()
        except Exception as e:  # pragma: no cover
            log("clock sampler unavailable:", e)

    def _pump(self):
        for line in self.proc.stdout:
            self.rows.append((time.time(), [c.strip() for c in line.split(",")]))

    def stop(self, t0, t1):
        if self.proc is None:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        time.sleep(0.15)
        self.proc.terminate()
        rows = [r for t, r in self.rows if t0 <= t <= t1 + 0.2] or [r for _, r in self.rows[-3:]]
        if not rows:
            return {"sm_mhz": None, "sm_max_mhz": None, "reasons": ["unsampled"]}
        sm = sorted(float(r[1]) for r in rows)
        names = ["hw_slowdown", "hw_thermal_slowdown", "sw_thermal_slowdown", "sw_power_cap"]
        reasons = sorted({nm for r in rows for nm, v in zip(names, r[5:9]) if v.lower().startswith("active")})
        return {"sm_mhz": sm[len(sm) // 2], "sm_max_mhz": float(rows[0][2]), "reasons": reasons,
                "power_w_max": max(float(r[3]) for r in rows), "samples": len(rows)}


def cpu_baseline(n_sample, eps, seconds_cap=30.0):
    """The reference's CPU force pass (oracle/_ref when built, else the oracle port) on a bounded
    sample of the same workload: the first n_sample bodies of the Plummer sphere (2-D projection,
    the reference is 2-D), one compute_gravity call, 1 thread (the reference is single-threaded)."""
    from oracle import bindings as ob
    import __graft_entry__ as ge
    pkg = ge.load_package()
    pos, vel, mass = pkg.scenarios.plummer(1 << 20, seed=1234, dim=2)
    pos, vel, mass = pos[:n_sample].copy(), vel[:n_sample].copy(), mass[:n_sample].copy()
    if ob.have_reference():
        kind = "reference"
        r = ob.Reference()
        r.set_config(softening=eps)
        r.load_bodies(pos, vel, mass)
        t0 = time.perf_counter()
        r.compute_gravity()
        dt = time.perf_counter() - t0
        r.close()
    else:
        kind = "port"
        o = ob.Oracle()
        t0 = time.perf_counter()
        o.compute_gravity(pos, vel, mass, softening=eps)
        dt = time.perf_counter() - t0
    inter = float(n_sample) * (n_sample - 1)   # N(N-1)/2 pair evaluations update both bodies
    return {"value": inter / dt, "unit": UNIT, "cores": 1, "kind": kind, "seconds": dt,
            "host_cores_available": os.cpu_count(),
            "sample": f"one compute_gravity pass (Physics.hpp:231-258, direct branch) over the first {n_sample} bodies "
                      f"of the 2^20 Plummer sphere projected to 2-D; N(N-1) interactions counted"}


def peak_fp32_tflops(sm_count, mhz):
    return sm_count * FP32_LANES_PER_SM * 2 * mhz * 1e6 / 1e12


def timed_steps(ph, dt, steps, barrier, reduce_max, diagnostics=False):
    """K back-to-back steps under ONE CUDA-event pair on the launch stream: no host barrier inside, so a
    sharded run's exchange overlaps the next step's own-sources launch exactly as in production; each step is
    preceded by a 256 MiB L2 flush on the same stream (its ~0.05 ms are inside the timed region) and
    nb_timer_end waits for the last exchange before it records. Returns the max over ranks in ms."""
    barrier()
    ph.timer_begin()
    for _ in range(steps):
        ph.flush_l2()
        ph.step(dt, 1, diagnostics=diagnostics)
    ms = ph.timer_end()
    barrier()
    return reduce_max(ms)


def fixture_parity(pkg, ph, dist, n):
    """Reference parity at this rank count, inside the driver-run bench: one compute_gravity over the 2^20-body
    sphere in the reference's 2-D layout, every rank's rows gathered, against 8192 accelerations the UNMODIFIED
    reference computed at full size (tests/golden/c4_plummer2d_2e20_subsample.npz, Physics.hpp:231-258).
    Bound: 1e-5 per body, scaled by the condition number of the sum where a body's pulls cancel (the rule of
    tests/test_gpu_fullsize.py::test_c4_full_size_against_the_reference_itself)."""
    path = os.path.join(ROOT, "tests", "golden", "c4_plummer2d_2e20_subsample.npz")
    if n != (1 << 20) or not os.path.exists(path):
        return None
    g = np.load(path)
    pos, vel, mass = pkg.scenarios.plummer(n, seed=1234, dim=2)
    eps = float(g["softening"])
    ph.set_config(softening=eps, integrator=1)
    ph.upload(pos, vel, mass)
    ph.compute_gravity()
    acc = ph.download()[2]
    if dist is not None:
        acc = pkg.sharded.gather_rows(dist, acc, n)
    idx = g["idx"]
    err = np.linalg.norm(acc[idx] - g["acc"], axis=1) / np.linalg.norm(g["acc"], axis=1)
    m64 = mass.astype(np.float64)
    worst, ok = 0.0, True
    for k in range(len(idx)):
        e = float(err[k])
        if e > 5e-6:  # evaluate the conditioning only where it can matter
            d = pos - pos[idx[k]]
            r2 = (d * d).sum(axis=1) + eps * eps
            t = (m64 / (r2 * np.sqrt(r2)))[:, None] * d
            cond = np.abs(t).sum() / np.linalg.norm(t.sum(axis=0))
            if cond > 32:
                ok = ok and (e / cond <= 6e-7)
                e = e * 32.0 / cond
        worst = max(worst, e)
    ok = bool(ok and worst <= 1e-5 and np.isfinite(acc).all())
    return {"fixture": "tests/golden/c4_plummer2d_2e20_subsample.npz (unmodified reference, full-size pass)",
            "max_rel_err": worst, "max_rel_err_raw": float(err.max()), "p999": float(np.quantile(err, 0.999)),
            "median": float(np.median(err)), "n_checked": int(len(idx)), "tolerance": 1e-5,
            "rule": "per-body |da|/|a| <= 1e-5; bodies with condition number c > 32 are scaled by 32/c", "ok": ok}


def state_digest(pkg, ph, dist, n):
    """Order-independent summary of the state after the timed steps, for comparison across rank counts."""
    pos, vel, _ = ph.download()
    if dist is not None:
        pos, vel = pkg.sharded.gather_rows(dist, pos, n), pkg.sharded.gather_rows(dist, vel, n)
    t0 = time.perf_counter()
    d = ph.compute_diagnostics()
    ph.synchronize()
    diag_s = time.perf_counter() - t0
    return {"sum_abs_x": float(np.abs(pos).sum()), "sum_abs_v": float(np.abs(vel).sum()), "energy": d.energy,
            "kinetic": d.kinetic, "potential": d.potential, "momentum": [float(v) for v in d.momentum],
            "ok": bool(d.ok),
            "agreement_expected": "sum_abs_x, sum_abs_v <= 1e-11 relative across rank counts (FP64 state, same "
                                  "arithmetic, different summation grouping of the sources); energy <= 1e-9"}, diag_s


def side_figure(pkg, device, pos, vel, mass, eps, dt, steps, env=None):
    """The same workload through another kernel variant, timed like the headline: warm-up, then `steps`
    back-to-back steps with an L2 flush before each, one CUDA-event pair, clocks sampled."""
    for k, v in (env or {}).items():
        os.environ[k] = v
    p = pkg.Physics(device)
    for k in (env or {}):
        os.environ.pop(k)
    p.set_precision(pkg.PRECISION_FP32)
    p.set_config(softening=eps, integrator=1)
    p.upload(pos, vel, mass)
    p.step(dt, 3)
    p.flush_l2()  # allocates the flush buffer outside the timed region
    p.synchronize()
    sampler = ClockSampler(device)
    sampler.start()
    t0 = time.time()
    ms = timed_steps(p, dt, steps, p.synchronize, lambda x: x)
    clocks = sampler.stop(t0, time.time())
    p.close()
    n = len(mass)
    return {"value": float(n) * n * steps / (ms * 1e-3), "unit": UNIT, "steps": steps, "ms_per_step": ms / steps,
            "l2": "flushed before every step", "clocks": clocks}

def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        log(f"note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE")
    import __graft_entry__ as ge
    pkg = ge.load_package()
    if args.config == 5:
        n = args.n if args.n_given else (1 << 22)
        pos, vel, mass = pkg.scenarios.two_galaxies(n, seed=1234, dim=3)
        eps = 3.844e8 / np.sqrt(n / 2)
        dt = 1e-3 * pkg.scenarios.crossing_time(n // 2)
        what = (f"config 5: {n}-body two-galaxy collision (two 3-D Plummer spheres on a bound orbit, equal masses, seed 1234), "
                f"Velocity Verlet, softening a/sqrt(N/2), dt 1e-3 crossing times, targets partitioned over {world} GPU(s)")
    else:
        n = args.n
        pkg, pos, vel, mass, eps, dt = workload(n, 3)
        what = (f"config 4: {n}-body Plummer sphere (3-D, equal masses, seed 1234), Velocity Verlet, "
                f"softening a/sqrt(N), dt 1e-3 crossing times, targets partitioned over {world} GPU(s)")
    dist = None
    if world > 1:
        import torch
        import torch.distributed as dist
        torch.cuda.set_device(local_rank)
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local_rank))
        ph = pkg.sharded.create(dist, local_rank)
    else:
        ph = pkg.Physics(local_rank)

    def barrier():
        ph.synchronize()
        if dist is not None:
            dist.barrier()

    def reduce_max(x):
        if dist is None:
            return x
        import torch
        t = torch.tensor([x], dtype=torch.float64, device="cuda")
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        return float(t.item())

    warm = max(args.warmup, 3)
    ph.set_precision(pkg.PRECISION_FP32)
    ph.set_config(softening=eps, integrator=1)
    ph.upload(pos, vel, mass)
    ph.step(dt, 1, diagnostics=True)  # first frame evaluates a(x0) like the gravity system, then one step
    e_first = ph.step_diagnostics().energy
    for _ in range(warm - 1):
        ph.flush_l2()  # also allocates the 256 MiB flush buffer here, outside the timed region
        ph.step(dt, 1)
    barrier()

    # ---- device-resident arm: K back-to-back steps, one CUDA-event pair on the launch stream
    sampler = ClockSampler(local_rank)
    c0 = ph.counters()
    ph.profile_forces(True)
    ph.overlap_profile(True)
    if rank == 0:
        sampler.start()
    t_wall0 = time.time()
    total_ms = timed_steps(ph, dt, args.steps, barrier, reduce_max)
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    n_force, force_ms = ph.profile_read()
    ph.profile_forces(False)
    overlap = ph.overlap_profile(False) if world > 1 else None
    c1 = ph.counters()
    value = float(n) * n * args.steps / (total_ms * 1e-3)

    # ---- what the timed steps left behind: a digest comparable across rank counts, energy drift, and the
    #      stand-alone diagnostics pass (hot loop B) timed on the same state
    ph.step(dt, 1, diagnostics=True)
    e_last = ph.step_diagnostics().energy
    digest, diag_s = state_digest(pkg, ph, dist, n)
    digest["steps_taken"] = warm + args.steps + 1
    digest["rel_energy_drift_fused"] = (e_last - e_first) / abs(e_first)
    diag_rec = {"seconds": diag_s, "pair_terms_per_s": float(n) * n / diag_s, "n_gpus": world,
                "note": "nb_compute_diagnostics (Physics.hpp:117-177): extent + pack + potential-only kernel over all "
                        "N^2 ordered pairs + readback, wall clock with a device synchronise"}

    # ---- end-to-end arm: host buffers in, host buffers out, every step (the drop-in's frame)
    lo, hi = ph.local_range()
    h_pos, h_vel, h_acc = (pkg.pinned_array((n, 3), np.float64) for _ in range(3))
    h_mass = pkg.pinned_array((n,), np.float32)
    h_mass[:] = mass
    ph.download_into(h_pos, h_vel, h_acc)
    if dist is not None:  # every rank needs the full state to re-upload it
        for h in (h_pos, h_vel, h_acc):
            h[:] = pkg.sharded.gather_rows(dist, h, n)
    e2e_ms = 0.0
    e2e_steps = max(2, min(args.steps, 3))
    for k in range(e2e_steps + 1):
        barrier()
        t0 = time.perf_counter()
        ph.upload_state(h_pos, h_vel, h_acc, h_mass)      # H2D of the frame's inputs
        ph.step(dt, 1)
        ph.download_into(h_pos, h_vel, h_acc)             # D2H of the frame's result rows (synchronises)
        if k > 0:  # first one warms the staging buffers
            e2e_ms += reduce_max((time.perf_counter() - t0) * 1e3)
        if dist is not None:  # untimed host plumbing: every rank re-uploads the full state next frame
            for h in (h_pos, h_vel, h_acc):
                h[:] = pkg.sharded.gather_rows(dist, h, n)
    e2e_value = float(n) * n * e2e_steps / (e2e_ms * 1e-3)
    h2d = (hi - lo) * (3 * 3 * 8 + 4 + 0)   # nb_upload_state copies a rank's own rows only (the rest arrive as FP32 sources over NVLink)
    d2h = (hi - lo) * 3 * 3 * 8

    general = planar = config3 = None
    if world == 1 and args.config == 4 and args.sides:
        # the same workload through the general-mass kernel (the production run above uses the
        # uniform-mass variant because every body of config 4 has the same mass)
        general = side_figure(pkg, local_rank, pos, vel, mass, eps, dt, max(args.steps, 5), env={"NB_FORCE_GENERIC": "1"})
        general["note"] = "NB_FORCE_GENERIC=1: per-source masses multiplied in the kernel (12 instead of 11 FMA-pipe ops per pair)"
        # and the reference's own layout: the same sphere projected to z = 0 and uploaded as 2-D
        # (Position/Velocity are DVec2), which selects the planar kernel variant
        p2, v2, m2 = pkg.scenarios.plummer(n, seed=1234, dim=2)
        planar = side_figure(pkg, local_rank, p2, v2, m2, eps, dt, max(args.steps, 5))
        planar["note"] = "2-D reference layout (z = 0): 8 FMA-pipe ops per pair; counted as interactions, not as 20-FLOP work"
        # config 3: 65536-body Plummer sphere, Verlet, fused energy/momentum diagnostics read back every step
        n3 = 65536
        p3, v3, m3 = pkg.scenarios.plummer(n3, seed=1234, dim=3)
        eps3, dt3 = 3.844e8 / np.sqrt(n3), 1e-3 * pkg.scenarios.crossing_time(n3)
        q = pkg.Physics(local_rank)
        q.set_precision(pkg.PRECISION_FP32)
        q.set_config(softening=eps3, integrator=1)
        q.upload(p3, v3, m3)
        q.step(dt3, 5)
        q.synchronize()
        k3 = 50
        q.timer_begin()
        q.step(dt3, k3)
        plain_ms = q.timer_end() / k3
        t0 = time.perf_counter()
        for _ in range(k3):
            q.step(dt3, 1, diagnostics=True)
            d3 = q.step_diagnostics()      # blocking read-back of the 12 sums
        diag_ms = (time.perf_counter() - t0) * 1e3 / k3
        q.timer_begin()
        for _ in range(10):
            q.compute_diagnostics()
        alone_ms = q.timer_end() / 10
        q.close()
        config3 = {"n_bodies": n3, "plain_step_ms": plain_ms, "step_with_diagnostics_readback_ms": diag_ms,
                   "interactions_per_s_plain": float(n3) * n3 / (plain_ms * 1e-3),
                   "interactions_per_s_with_diagnostics": float(n3) * n3 / (diag_ms * 1e-3),
                   "standalone_diagnostics_ms": alone_ms, "energy": d3.energy,
                   "note": "config 3: per-step energy/momentum diagnostics from the fused potential (wall clock incl. read-back)"}

    tree = None
    if world == 1 and args.tree and args.config == 4:
        # The reference's DEFAULT solver at this body count: above bh_threshold = 100 bodies its
        # compute_gravity is a float Barnes-Hut quadtree (Physics.hpp:212-230), reproduced bit for bit
        # by NB_SOLVER_REFERENCE. Not the headline metric (a tree pass does ~600 cell visits per
        # body, not N interactions): reported as time per Verlet step beside the reference's own.
        p2, v2, m2 = pkg.scenarios.plummer(n, seed=1234, dim=2)
        pt = pkg.Physics(local_rank)
        pt.set_config(softening=eps, integrator=1)
        pt.set_solver(pkg.SOLVER_REFERENCE, pkg.REF_BH_THRESHOLD, pkg.REF_BH_THETA)
        pt.solver_profile(True)
        pt.upload(p2, v2, m2)
        pt.step(dt, 1)
        info = pt.solver_info()
        pt.solver_profile(False)
        pt.step(dt, 2)
        pt.synchronize()
        pt.timer_begin()
        pt.step(dt, 10)
        t_ms = pt.timer_end() / 10
        pt.close()
        tree = {"ms_per_step": t_ms, "steps_per_s": 1e3 / t_ms, "bodies_per_s": n * 1e3 / t_ms,
                "cells": info["cells"], "depth": info["depth"], "cells_visited_per_body": info["visited"] / n,
                "terms_per_body": info["terms"] / n, "cell_visits_per_s": info["visited"] * 1e3 / t_ms,
                "note": "NB_SOLVER_REFERENCE: the reference's default Barnes-Hut branch (bh_threshold 100, theta 0.5), "
                        "bit-identical accelerations; 2-D projection of the same sphere; one tree build + walk per step"}
        if rank == 0 and args.cpu_sample > 0:
            from oracle import bindings as ob
            if ob.have_reference():
                r = ob.Reference()
                r.set_config(softening=eps)
                r.set_solver(ob.REF_BH_THRESHOLD, ob.REF_BH_THETA)
                r.load_bodies(p2, v2, m2)
                t0 = time.perf_counter()
                r.compute_gravity()
                tree["cpu_seconds_per_pass"] = time.perf_counter() - t0
                tree["cpu_kind"] = "reference"
                r.close()
            else:
                t0 = time.perf_counter()
                ob.Oracle().bh_gravity(p2, v2, m2, softening=eps)
                tree["cpu_seconds_per_pass"] = time.perf_counter() - t0
                tree["cpu_kind"] = "port"

    # ---- reference parity at THIS rank count (every rank takes part; rank 0 reports)
    near = None
    parity = None
    if args.config == 4 and args.parity:
        ph.near_stats(True)
        parity = fixture_parity(pkg, ph, dist, n)
        visits, near_visits = ph.near_stats(False)
        near = {"warp_group_visits": visits, "near_path_visits": near_visits,
                "near_fraction": near_visits / max(visits, 1),
                "note": "share of (warp, 128-row source group) visits that took the two-float path in the parity pass"}

    status = 0
    if rank == 0:
        props_sms = ph.device_info()["sm_count"]
        peaks = {}
        try:
            peaks = json.load(open(os.path.join(ROOT, "MEASURED_PEAKS.json")))
        except Exception:
            pass
        max_mhz = float(peaks.get("sm_max_mhz", 1965.0))
        peak = peak_fp32_tflops(props_sms, max_mhz)
        per_gpu_inter = float(n) * (hi - lo)   # rank 0's share of one force evaluation
        kern_s = force_ms * 1e-3 / max(args.steps, 1)  # force-kernel time per step (own + remote launches)
        achieved = per_gpu_inter * FLOPS_PER_INTERACTION / kern_s / 1e12
        traffic = None
        try:
            traffic = json.load(open(os.path.join(ROOT, "profiles", "roofline_traffic.json")))["dram_bytes_per_launch"]
        except Exception:
            pass
        roof = {"bound": "fp32", "achieved": achieved, "peak": peak, "unit": "TFLOP/s", "frac": achieved / peak,
                "traffic": traffic,
                "peak_source": f"{props_sms} SMs x 128 FP32 lanes x 2 x sm_max_mhz {max_mhz:.0f} (MEASURED_PEAKS.json clock; "
                               "the file has no FP32 CUDA-core figure, so the peak is derived from its clock)",
                "flops_per_interaction": FLOPS_PER_INTERACTION, "force_launches": n_force,
                "force_ms_per_step": kern_s * 1e3}
        if clocks and clocks.get("sm_mhz"):
            roof["frac_at_observed_clock"] = achieved / peak_fp32_tflops(props_sms, clocks["sm_mhz"])
        n_cpu = args.cpu_sample
        cpu = cpu_baseline(n_cpu, eps) if n_cpu > 0 else None
        line = {
            "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": world, "steps": args.steps,
            "warmup": warm, "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 pair terms / f64 accumulation+state",
            "data": "synthetic",
            "config": {"workload": what, "n_bodies": n, "interactions_per_step": float(n) * n,
                       "integrator": "velocity_verlet",
                       "l2": "flushed (256 MiB memset on the launch stream) before every timed step",
                       "timing": "one CUDA-event pair around the K back-to-back steps (flushes included, no host barrier "
                                 "inside; the pair closes after the last step's exchange), max over ranks",
                       "parallelism": f"target-partition x{world}"},
            "pct_fp32_peak": 100.0 * value * FLOPS_PER_INTERACTION / (peak * 1e12 * world),
            "steps_per_s": 1e3 * args.steps / total_ms,
            "kernel_variant": "uniform-mass (all bodies of this workload have equal mass)", "general_mass": general,
            "planar_2d": planar, "config3": config3, "reference_solver": tree, "diagnostics": diag_rec,
            "roofline": roof, "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
                    "path": "nb_upload_state (pinned host) -> nb_step -> nb_download3 (pinned host), wall clock"},
            "gpu_launches": c1["kernel_launches"] - c0["kernel_launches"],
            "launches_note": "kernels of this library inside the timed region, per rank: one fused force launch per step at "
                             "1 GPU, two (own / remote sources) when sharded; the L2-flush memsets are not kernels of the library",
            "parity": parity, "near_path": near, "state_digest": digest, "exchange_overlap": overlap,
            "clocks": clocks,
        }
        emit(line)
        if parity is not None and not parity["ok"]:
            log(f"PARITY FAILED at {world} GPU(s): {parity}")
            status = 3
        if not digest["ok"]:
            log("diagnostics of the end state are not finite")
            status = 3
    ph.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()
    return status


def run_reference(args):
    """The reference's own CPU implementation of the path, timed on this box's host cores."""
    rank = int(os.environ.get("RANK", "0"))
    if rank != 0:
        return
    n = args.n
    eps = 3.844e8 / np.sqrt(n)
    steps = max(1, args.steps)
    ns = args.ref_sample
    best = None
    vals = []
    for k in range(max(args.warmup, 1) + steps):
        cb = cpu_baseline(ns, eps)
        if k >= max(args.warmup, 1):
            vals.append(cb)
    total_s = sum(v["seconds"] for v in vals)
    value = float(ns) * (ns - 1) * len(vals) / total_s
    best = vals[0]
    line = {
        "impl": "reference", "metric": METRIC, "value": value, "unit": UNIT, "n_gpus": args.gpus, "steps": steps,
        "warmup": max(args.warmup, 1), "ms_per_step": 1e3 * total_s / len(vals), "higher_is_better": True,
        "scaling": "strong", "vs_baseline": None, "dtype": "f64", "data": "synthetic",
        "config": {"workload": f"config 4: {n}-body Plummer sphere, bounded sample of {ns} bodies per step "
                               "(the reference's O(N^2) single-threaded loop needs ~1 h for one full pass)",
                   "n_bodies": n, "sample_bodies": ns},
        "cpu_baseline": {"value": value, "unit": UNIT, "cores": 1, "kind": best["kind"], "sample": best["sample"],
                         "host_cores_available": os.cpu_count(),
                         "note": "the reference is single-threaded (no OpenMP/threads, flecs workers never enabled)"},
        "e2e": {"value": value, "unit": UNIT, "h2d_bytes_per_step": 0, "d2h_bytes_per_step": 0},
    }
    emit(line)


def main():
    # libraries (NCCL's version banner, for one) may print to fd 1: keep the real stdout for the one
    # JSON line and send everything else to stderr
    global _JSON_OUT
    _JSON_OUT = os.fdopen(os.dup(1), "w")
    os.dup2(2, 1)
    ap = argparse.ArgumentParser()
    ap.add_argument("--gpus", type=int, default=1)
    ap.add_argument("--steps", type=int, default=5)
    ap.add_argument("--warmup", type=int, default=3)
    ap.add_argument("--impl", default="ours", choices=["ours", "reference"])
    ap.add_argument("--n", type=int, default=None, help="bodies (BASELINE metric: 2^20; config 5: 2^22)")
    ap.add_argument("--config", type=int, default=4, choices=[4, 5], help="BASELINE.json configs[3] (default) or configs[4]")
    ap.add_argument("--no-parity", dest="parity", action="store_false", help="skip the fixture parity pass")
    ap.add_argument("--no-sides", dest="sides", action="store_false",
                    help="skip the general-mass / planar / config-3 side figures")
    ap.add_argument("--no-tree", dest="tree", action="store_false",
                    help="skip the extra figure for the reference's default Barnes-Hut solver")
    # sample sizes are deliberately not powers of two: the reference's loop streams pos[j] and acc[j]
    # (both 16 B/body); at N = 16384 the two arrays sit 256 KiB apart and alias in the host caches,
    # which made that size 2-4x slower per interaction than its neighbours (profiles/r01_cpu_baseline_table.txt)
    ap.add_argument("--cpu-sample", type=int, default=30000, help="bodies in the cpu_baseline sample (0 = skip)")
    ap.add_argument("--ref-sample", type=int, default=20000, help="bodies per step of --impl reference")
    args = ap.parse_args()
    args.n_given = args.n is not None
    if args.n is None:
        args.n = 1 << 20
    if args.impl == "reference":
        run_reference(args)
        return 0
    return run_ours(args)


if __name__ == "__main__":
    sys.exit(main())
