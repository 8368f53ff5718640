"""Developer timing on a B200: ms per fused Verlet step (L2 flushed, back to back) and the share of
(warp, source group) visits that took the near path, for a list of body counts.
    python tools/time_step.py 1048576:3 65536:50 [plummer|disk] [dim]"""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
sc = pkg.scenarios
args = [a for a in sys.argv[1:] if ":" in a] or ["1048576:3"]
shape = "disk" if "disk" in sys.argv else "plummer"
dim = 2 if "2" in sys.argv[1:] else 3
for a in args:
    n, steps = (int(v) for v in a.split(":"))
    if shape == "disk":
        pos, vel, mass = sc.uniform_disk(n, seed=1234, circular=True)
    else:
        pos, vel, mass = sc.plummer(n, seed=1234, dim=dim)
    eps, dt = 3.844e8 / np.sqrt(n), 1e-3 * sc.crossing_time(n)
    ph = pkg.Physics(0)
    ph.set_precision(pkg.PRECISION_FP32)
    ph.set_config(softening=eps, integrator=1)
    ph.upload(pos, vel, mass)
    ph.step(dt, 3)
    ph.synchronize()
    k0 = ph.counters()["kernel_launches"]
    ph.timer_begin()
    for _ in range(steps):
        ph.flush_l2()
        ph.step(dt, 1)
    ms = ph.timer_end() / steps
    k1 = ph.counters()["kernel_launches"]
    v = nv = 0
    if hasattr(ph._L, "nb_near_stats"):
        ph.near_stats(True)
        ph.step(dt, 1)
        v, nv = ph.near_stats(False)
    ph.timer_begin()
    for _ in range(max(2, steps // 4)):
        ph.compute_diagnostics()
    dms = ph.timer_end() / max(2, steps // 4)
    ips = float(n) * n / (ms * 1e-3)
    print(f"{os.environ.get('NB_LIB_PATH', 'default')} {shape}{dim}d n={n}: {ms:.4f} ms/step {ips:.4e} inter/s "
          f"({ips * 20 / 74.45e12 * 100:.2f}% of 74.45 TF) launches/step {(k1 - k0) / steps:.2f} near {nv}/{v} = {nv / max(v, 1):.5f} "
          f"| diagnostics {dms:.4f} ms = {float(n) * n / (dms * 1e-3):.4e} pair terms/s", flush=True)
    ph.close()
