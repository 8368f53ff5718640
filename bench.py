#!/usr/bin/env python
"""bench.py — body-body interactions/s of the physics step on B200 (BASELINE.json metric).

    python bench.py --gpus N --steps K --warmup W            # this repo's CUDA path
    python bench.py --impl reference --steps K --warmup W    # the reference's CPU code, same metric

A step = one fused Velocity Verlet step (one force evaluation over N^2 body-body interactions
+ kick/drift/pack in its epilogue) of config 4: the 2^20-body Plummer sphere, target-partitioned
over the ranks (one process per GPU, NCCL all-gather of packed positions per step). One JSON
line on stdout from rank 0; everything else goes to stderr.
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


def run_ours(args):
    rank = int(os.environ.get("RANK", "0"))
    local_rank = int(os.environ.get("LOCAL_RANK", "0"))
    world = int(os.environ.get("WORLD_SIZE", "1"))
    if world != args.gpus:
        log(f"note: --gpus {args.gpus} but WORLD_SIZE={world}; using WORLD_SIZE")
    n = args.n
    pkg, pos, vel, mass, eps, dt = workload(n, 3)
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

    ph.set_precision(pkg.PRECISION_FP32)
    ph.set_config(softening=eps, integrator=1)
    ph.upload(pos, vel, mass)
    ph.step(dt, 1)  # first frame evaluates a(x0) like the gravity system, then one step
    for _ in range(max(args.warmup, 3) - 1):
        ph.step(dt, 1)
    barrier()

    # ---- device-resident arm: K steps, L2 flushed between steps, CUDA events on the launch stream
    sampler = ClockSampler(local_rank)
    c0 = ph.counters()
    ph.profile_forces(True)
    if rank == 0:
        sampler.start()
    t_wall0 = time.time()
    total_ms = 0.0
    for _ in range(args.steps):
        ph.flush_l2()
        barrier()
        ph.timer_begin()
        ph.step(dt, 1)
        ms = ph.timer_end()
        barrier()
        total_ms += reduce_max(ms)
    t_wall1 = time.time()
    clocks = sampler.stop(t_wall0, t_wall1) if rank == 0 else None
    n_force, force_ms = ph.profile_read()
    ph.profile_forces(False)
    c1 = ph.counters()
    value = float(n) * n * args.steps / (total_ms * 1e-3)

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
        barrier()
        if k > 0:  # first one warms the staging buffers
            e2e_ms += reduce_max((time.perf_counter() - t0) * 1e3)
        if dist is not None:  # untimed host plumbing: every rank re-uploads the full state next frame
            for h in (h_pos, h_vel, h_acc):
                h[:] = pkg.sharded.gather_rows(dist, h, n)
    e2e_value = float(n) * n * e2e_steps / (e2e_ms * 1e-3)
    h2d = (hi - lo) * (3 * 3 * 8 + 4 + 0)   # nb_upload_state copies a rank's own rows only (the rest arrive as FP32 sources over NVLink)
    d2h = (hi - lo) * 3 * 3 * 8

    general = None
    planar = None
    if world == 1:
        # the same workload through the general-mass kernel (the production run above uses the
        # uniform-mass variant because every body of config 4 has the same mass)
        os.environ["NB_FORCE_GENERIC"] = "1"
        pg = pkg.Physics(local_rank)
        os.environ.pop("NB_FORCE_GENERIC")
        pg.set_precision(pkg.PRECISION_FP32)
        pg.set_config(softening=eps, integrator=1)
        pg.upload(pos, vel, mass)
        pg.step(dt, 2)
        pg.synchronize()
        pg.timer_begin()
        pg.step(dt, 2)
        g_ms = pg.timer_end()
        pg.close()
        general = {"value": float(n) * n * 2 / (g_ms * 1e-3), "unit": UNIT, "steps": 2,
                   "note": "NB_FORCE_GENERIC=1: per-source masses multiplied in the kernel (12 instead of 11 FMA-pipe ops per pair)"}
        # and the reference's own layout: the same sphere projected to z = 0 and uploaded as 2-D
        # (Position/Velocity are DVec2), which selects the planar kernel variant
        p2, v2, m2 = pkg.scenarios.plummer(n, seed=1234, dim=2)
        pp = pkg.Physics(local_rank)
        pp.set_precision(pkg.PRECISION_FP32)
        pp.set_config(softening=eps, integrator=1)
        pp.upload(p2, v2, m2)
        pp.step(dt, 2)
        pp.synchronize()
        pp.timer_begin()
        pp.step(dt, 2)
        p_ms = pp.timer_end()
        pp.close()
        planar = {"value": float(n) * n * 2 / (p_ms * 1e-3), "unit": UNIT, "steps": 2,
                  "note": "2-D reference layout (z = 0): 8 FMA-pipe ops per pair; counted as interactions, not as 20-FLOP work"}

    tree = None
    if world == 1 and args.tree:
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
        kern_s = force_ms * 1e-3 / max(n_force, 1) * (n_force / max(args.steps, 1))  # force time per step
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
            "warmup": max(args.warmup, 3), "ms_per_step": total_ms / args.steps, "higher_is_better": True,
            "scaling": "strong", "vs_baseline": None, "dtype": "f32 pair terms / f64 accumulation+state",
            "data": "synthetic",
            "config": {"workload": f"config 4: {n}-body Plummer sphere (3-D, equal masses, seed 1234), Velocity Verlet, "
                                   f"softening a/sqrt(N), dt 1e-3 crossing times, targets partitioned over {world} GPU(s)",
                       "n_bodies": n, "interactions_per_step": float(n) * n, "integrator": "velocity_verlet",
                       "l2": "flushed (256 MiB memset) before every timed step", "parallelism": f"target-partition x{world}"},
            "pct_fp32_peak": 100.0 * value * FLOPS_PER_INTERACTION / (peak * 1e12 * world),
            "steps_per_s": 1e3 * args.steps / total_ms,
            "kernel_variant": "uniform-mass (all bodies of config 4 have equal mass)", "general_mass": general, "planar_2d": planar,
            "reference_solver": tree,
            "roofline": roof, "cpu_baseline": cpu,
            "e2e": {"value": e2e_value, "unit": UNIT, "h2d_bytes_per_step": h2d, "d2h_bytes_per_step": d2h,
                    "steps": e2e_steps, "ms_per_step": e2e_ms / e2e_steps,
                    "path": "nb_upload_state (pinned host) -> nb_step -> nb_download3 (pinned host), wall clock"},
            "gpu_launches": c1["kernel_launches"] - c0["kernel_launches"],
            "clocks": clocks,
        }
        emit(line)
    ph.close()
    if dist is not None:
        dist.barrier()
        dist.destroy_process_group()


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
    ap.add_argument("--n", type=int, default=1 << 20, help="bodies (BASELINE metric: 2^20)")
    ap.add_argument("--no-tree", dest="tree", action="store_false",
                    help="skip the extra figure for the reference's default Barnes-Hut solver")
    # sample sizes are deliberately not powers of two: the reference's loop streams pos[j] and acc[j]
    # (both 16 B/body); at N = 16384 the two arrays sit 256 KiB apart and alias in the host caches,
    # which made that size 2-4x slower per interaction than its neighbours (profiles/r01_cpu_baseline_table.txt)
    ap.add_argument("--cpu-sample", type=int, default=30000, help="bodies in the cpu_baseline sample (0 = skip)")
    ap.add_argument("--ref-sample", type=int, default=20000, help="bodies per step of --impl reference")
    args = ap.parse_args()
    if args.impl == "reference":
        run_reference(args)
    else:
        run_ours(args)


if __name__ == "__main__":
    main()
