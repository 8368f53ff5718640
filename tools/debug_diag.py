"""Developer check: stand-alone diagnostics on an offset cluster, piece by piece."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
for off in (0.0, 1.0e6, 1.0e9):
    for dim in (2, 3):
        for n in (2000, 300):
            pos, vel, mass = pkg.scenarios.plummer(n, seed=3, dim=dim, scale=1.0e3)
            pos = pos + off
            ph = pkg.Physics(0)
            ph.set_precision(pkg.PRECISION_FP32)
            ph.set_config(softening=1.0)
            ph.upload(pos, vel, mass)
            d = ph.compute_diagnostics()
            ph.step(1e-6, 1, diagnostics=True)
            d2 = ph.step_diagnostics()
            m = mass.astype(np.float64)
            pe = 0.0
            for i in range(n):
                dd = pos[i + 1:] - pos[i]
                pe -= 6.6743e-11 * m[i] * (m[i + 1:] / np.sqrt((dd * dd).sum(axis=1) + 1.0)).sum()
            print(f"off={off:g} dim={dim} n={n}: standalone PE {d.potential:.6e} ok={d.ok} fused PE {d2.potential:.6e} numpy {pe:.6e}", flush=True)
            ph.close()
