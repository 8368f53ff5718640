"""Developer timing of the BASELINE configs' shapes on one GPU -> stdout.
  planar : 2^20-body Plummer projected to z = 0 (reference 2-D layout), fused Verlet step
  c3     : 65536-body Plummer, Verlet, with and without per-step fused diagnostics
  small  : frame rate of nb_step(dt, 1) for reference-sized worlds (launch/latency bound)"""
import os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package(); sc = pkg.scenarios
what = sys.argv[1:] or ["planar", "c3", "small", "overlap"]

def timed(ph, fn, reps=3):
    best = 1e30
    for _ in range(reps):
        ph.timer_begin(); fn(); best = min(best, ph.timer_end())
    return best

if "planar" in what:
    n = 1 << 20
    for dim in (2, 3):
        pos, vel, mass = sc.plummer(n, dim=dim); dt = 1e-3 * sc.crossing_time(n); eps = 3.844e8 / np.sqrt(n)
        for generic in ("0", "1"):
            os.environ["NB_FORCE_GENERIC"] = generic
            ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.set_config(softening=eps)
            ph.upload(pos, vel, mass); ph.step(dt, 2); ph.synchronize()
            ms = timed(ph, lambda: ph.step(dt, 1))
            print(f"planar: n={n} dim={dim} generic={generic}: {ms:.3f} ms/step {n*float(n)/ms/1e9:.4f}e12 inter/s "
                  f"({n*float(n)/ms/1e9/3.72250*100:.1f}% of 20-flop peak)", flush=True)
            ph.close()
    os.environ.pop("NB_FORCE_GENERIC")
if "c3" in what:
    n = 65536
    pos, vel, mass = sc.plummer(n, dim=3); dt = 1e-3 * sc.crossing_time(n); eps = 3.844e8 / np.sqrt(n)
    ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.set_config(softening=eps)
    ph.upload(pos, vel, mass); ph.step(dt, 2); ph.synchronize()
    a = timed(ph, lambda: ph.step(dt, 20)) / 20
    def with_diag():
        for _ in range(20):
            ph.step(dt, 1, diagnostics=True); ph.step_diagnostics()
    b = timed(ph, with_diag) / 20
    print(f"c3: n={n}: step {a:.4f} ms ({n*float(n)/a/1e9:.4f}e12 inter/s), step+fused diagnostics read back {b:.4f} ms "
          f"({n*float(n)/b/1e9:.4f}e12 inter/s)", flush=True)
    ph.close()
if "small" in what:
    for n in (3, 256, 1024, 4096, 16384):
        pos, vel, mass = sc.uniform_disk(n, seed=3, circular=True)
        for prec, name in ((pkg.PRECISION_AUTO, "auto"), (pkg.PRECISION_FP32, "fp32")):
            ph = pkg.Physics(0); ph.set_precision(prec); ph.set_config(softening=1e6)
            ph.upload(pos, vel, mass); ph.step(100.0, 3); ph.synchronize()
            t0 = time.perf_counter()
            for _ in range(200): ph.step(100.0, 1)
            ph.synchronize(); t1 = time.perf_counter()
            t2 = time.perf_counter(); ph.step(100.0, 200); ph.synchronize(); t3 = time.perf_counter()
            print(f"small: n={n} {name}: {(t1-t0)/200*1e6:.1f} us/frame one call per frame, {(t3-t2)/200*1e6:.1f} us/frame batched", flush=True)
            ph.close()
if "overlap" in what:
    for n in (65536, 1 << 20):
        pos, vel, mass = sc.plummer(n, dim=2)
        ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.upload(pos, vel, mass)
        rad = np.full(n, 1.0e3)
        ph.count_overlaps(rad)
        t0 = time.perf_counter(); got = ph.count_overlaps(rad); t1 = time.perf_counter()
        print(f"overlap: n={n}: {1e3*(t1-t0):.2f} ms incl. radius upload, {n*(n-1)/2/(t1-t0):.3e} pairs/s, overlaps {got[0]}", flush=True)
        ph.close()
if "diag" in what:
    for n in (65536, 1 << 20):
        pos, vel, mass = sc.plummer(n, dim=3); eps = 3.844e8 / np.sqrt(n); dt = 1e-3 * sc.crossing_time(n)
        ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.set_config(softening=eps); ph.upload(pos, vel, mass)
        ph.compute_diagnostics()
        t0 = time.perf_counter(); d = ph.compute_diagnostics(); t1 = time.perf_counter()
        ph.step(dt, 1, diagnostics=True); ph.synchronize()
        t2 = time.perf_counter(); ph.step(dt, 1, diagnostics=True); d2 = ph.step_diagnostics(); t3 = time.perf_counter()
        print(f"diag: n={n}: stand-alone compute_diagnostics {1e3*(t1-t0):.2f} ms; step with fused diagnostics {1e3*(t3-t2):.2f} ms", flush=True)
        ph.close()
