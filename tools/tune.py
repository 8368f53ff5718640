"""Developer sweep of force-kernel variants at one N -> stdout (interactions/s)."""
import os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package(); sc = pkg.scenarios
n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 18
steps = int(sys.argv[2]) if len(sys.argv) > 2 else 4
variants = [tuple(int(x) for x in v.split(",")) for v in sys.argv[3:]] or [(4,64,4),(4,64,5),(3,64,4),(3,64,5),(3,64,6),(2,64,6),(2,64,8),(4,64,3)]
pos, vel, mass = sc.plummer(n, dim=3); dt = 1e-3 * sc.crossing_time(n); eps = 3.844e8 / np.sqrt(n)
for r, ch, mb in variants:
    os.environ.update(NB_FORCE_R=str(r), NB_FORCE_CHUNK=str(ch), NB_FORCE_MINB=str(mb))
    ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.set_config(softening=eps)
    ph.upload(pos, vel, mass); ph.step(dt, 2); ph.synchronize()
    best = 1e30
    for _ in range(3):
        ph.timer_begin(); ph.step(dt, steps); best = min(best, ph.timer_end())
    ips = n * float(n) * steps / (best * 1e-3)
    print(f"n={n} R={r} chunk={ch} minb={mb}: {best/steps:.3f} ms/step {ips:.4e} inter/s {ips*20/74.45e12*100:.2f}%", flush=True)
    ph.close()
