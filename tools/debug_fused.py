import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
from oracle import bindings as ob
o = ob.Oracle()
def run(n, integ, nseg, r, steps=1, dt=300.0, eps=1.0e6):
    os.environ["NB_FORCE_NSEG"] = str(nseg); os.environ["NB_FORCE_R"] = str(r)
    pos, vel, mass = pkg.scenarios.uniform_disk(n, seed=1234, circular=True)
    ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.set_config(integrator=integ, softening=eps)
    ph.upload(pos, vel, mass); ph.step(dt, steps); p, v, a = ph.download(); ph.close()
    po, vo = pos.copy(), vel.copy(); ao = np.zeros_like(pos); a0 = np.zeros_like(pos)
    o.physics_frames(po, vo, ao, a0, mass, np.zeros(n, np.uint8), ob.make_config(integrator=integ, softening=eps), dt, steps)
    ex = np.abs(p - po).max(axis=1) / 3.844e8; ev = np.abs(v - vo).max(axis=1); ea = np.abs(a-ao).max(axis=1)/np.abs(ao).max()
    bad = np.nonzero(ex > 1e-9)[0]
    print(f"eps={eps:g} n={n} integ={integ} nseg={nseg} R={r} steps={steps}: max ex {ex.max():.2e} ev {ev.max():.2e} ea {ea.max():.2e} nbad {bad.size} first bad {bad[:8]} last {bad[-3:]}", flush=True)
for n in (2048, 4096):
    for integ in (0, 1):
        for nseg, r in ((1, 1), (3, 4)):
            for steps in (1, 2, 5, 20):
                for eps in (1.0e6, 1.0e5):
                    run(n, integ, nseg, r, steps, 300.0 if n == 4096 else 600.0, eps)
