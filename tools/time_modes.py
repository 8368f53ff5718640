"""Developer timing: force-only pass vs fused Verlet step vs Euler step, with NB_FORCE_NSEG variants -> stdout."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package(); sc = pkg.scenarios
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
pos, vel, mass = sc.plummer(n, dim=3); dt = 1e-3 * sc.crossing_time(n); eps = 3.844e8 / np.sqrt(n)
for nseg in (sys.argv[2:] or ["0"]):
    os.environ["NB_FORCE_NSEG"] = nseg
    for integ in (1, 0):
        ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.set_config(softening=eps, integrator=integ)
        ph.upload(pos, vel, mass); ph.step(dt, 2); ph.synchronize()
        res = {}
        for name, fn in (("force_only", lambda: ph.compute_gravity()), ("step", lambda: ph.step(dt, 1))):
            best = 1e30
            for _ in range(3):
                ph.timer_begin(); fn(); best = min(best, ph.timer_end())
            res[name] = best
        print(f"n={n} nseg={nseg} integ={integ}: force_only {res['force_only']:.3f} ms  step {res['step']:.3f} ms  "
              f"-> {n*float(n)/res['force_only']/1e9:.4f}e12 / {n*float(n)/res['step']/1e9:.4f}e12 inter/s", flush=True)
        ph.close()
