"""Developer script for ncu captures: one POT step pair and two stand-alone diagnostics passes at N = 2^20 (3-D)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__)))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 20
pos, vel, mass = pkg.scenarios.plummer(n, seed=1234, dim=3)
eps, dt = 3.844e8 / np.sqrt(n), 1e-3 * pkg.scenarios.crossing_time(n)
ph = pkg.Physics(0)
ph.set_precision(pkg.PRECISION_FP32)
ph.set_config(softening=eps, integrator=1)
ph.upload(pos, vel, mass)
ph.step(dt, 2, diagnostics=True)      # launches: force-only (a(x0)), POT, POT
d = ph.step_diagnostics()
d2 = ph.compute_diagnostics()         # PONLY
d3 = ph.compute_diagnostics()
print("fused E", d.energy, "stand-alone E", d2.energy, d3.energy)
ph.close()
