"""Developer check on a B200: parity numbers and force-kernel variant timings -> stdout."""
import json, os, sys, time
ROOT = os.path.dirname(os.path.dirname(os.path.dirname(os.path.abspath(__file__))))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
pkg = ge.load_package()
from oracle import bindings as ob
sc = pkg.scenarios
o = ob.Oracle()

def relerr(a, ref):
    return np.linalg.norm(a - ref, axis=1) / np.linalg.norm(ref, axis=1)

def accuracy(n=4096):
    pos, vel, mass = sc.uniform_disk(n)
    ref = o.compute_gravity(pos, vel, mass)
    for r in (4, 3, 2, 1):
        os.environ["NB_FORCE_R"] = str(r)
        ph = pkg.Physics(0); ph.set_precision(pkg.PRECISION_FP32); ph.upload(pos, vel, mass); ph.compute_gravity(); _, _, acc = ph.download()
        e = relerr(acc, ref)
        print(f"C2 acc n={n} R={r}: max {e.max():.3e} p99.9 {np.quantile(e,0.999):.3e} med {np.median(e):.3e}", flush=True)
        ph.close()
    os.environ.pop("NB_FORCE_R")

def three_body():
    dt = pkg.dt_eff(np.float32(1.0)/np.float32(120), 1e6)
    for packed in (2, 1):
        for integ in (1, 0):
            ph = pkg.Physics(0); ph.set_config(integrator=integ)
            ph.set_precision(pkg.PRECISION_AUTO if packed == 2 else pkg.PRECISION_FP32)
            ph.reset_scenario()
            po, vo, mo, pino = o.seed_three_body()
            acc = np.zeros((3,2)); ap = np.zeros((3,2))
            cfg = ob.make_config(integrator=integ)
            d0 = ph.compute_diagnostics(); emax = 0.0; done = 0; odone = 0
            for upto in (100, 1000, 10000):
                while done < upto:
                    ph.step(dt, 100); done += 100
                    d = ph.compute_diagnostics(); emax = max(emax, abs((d.energy-d0.energy)/d0.energy))
                o.physics_frames(po, vo, acc, ap, mo, pino, cfg, dt, upto - odone); odone = upto
                p, v, a = ph.download()
                print(f"C1 mode={["fp32-scalar","fp32-packed","exact"][packed]} integ={integ} steps={upto}: bitexact={np.array_equal(p,po) and np.array_equal(v,vo)} max|dx|/R {np.abs(p-po).max()/3.844e8:.3e} max|dE/E0| {emax:.3e} body0 {p[0]} P {d.momentum[:2]}", flush=True)
            ph.close()

def timing(n, steps=3):
    pos, vel, mass = sc.plummer(n, dim=3)
    dt = 1e-3 * sc.crossing_time(n)
    for packed in (1,):
        for r in (4, 3, 2, 1):
            for nseg in (0,):
                os.environ["NB_FORCE_R"] = str(r); os.environ["NB_FORCE_NSEG"] = str(nseg)
                ph = pkg.Physics(0); ph.upload(pos, vel, mass)
                ph.step(dt, 1); ph.synchronize()
                ph.timer_begin(); ph.step(dt, steps); ms = ph.timer_end()
                ips = n * float(n) * steps / (ms * 1e-3)
                print(f"T n={n} packed={packed} R={r} nseg={nseg}: {ms/steps:.3f} ms/step  {ips:.3e} inter/s  {ips*20/74.4e12*100:.1f}% of 74.4TF", flush=True)
                ph.close()

if __name__ == "__main__":
    what = sys.argv[1:] or ["acc", "c1", "t16", "t20"]
    if "acc" in what: accuracy()
    if "c1" in what: three_body()
    if "t16" in what: timing(65536, 10)
    if "t18" in what: timing(262144, 4)
    if "t20" in what: timing(1 << 20, 2)
