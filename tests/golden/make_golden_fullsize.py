"""Full-size golden vector from the UNMODIFIED reference (oracle/_ref/libnbref.so): one
compute_gravity pass (Physics.hpp:183-261, direct branch) over the 2^20-body Plummer sphere of
config 4 projected to 2-D (the reference is 2-D), softening a/sqrt(N). About half an hour to an hour
on one core, so it is generated once here and committed as a strided subsample:
  tests/golden/c4_plummer2d_2e20_subsample.npz   idx (every 128th body), acc[idx], sum_i m_i a_i
Run:  python tests/golden/make_golden_fullsize.py"""
import os, sys, time
HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge
from oracle import bindings as ob

def main():
    pkg = ge.load_package()
    n = 1 << 20
    pos, vel, mass = pkg.scenarios.plummer(n, seed=1234, dim=2)
    eps = 3.844e8 / np.sqrt(n)
    r = ob.Reference()
    r.set_config(softening=eps)
    r.load_bodies(pos, vel, mass)
    t0 = time.time()
    r.compute_gravity()
    dt = time.time() - t0
    acc = r.read_bodies()[2]
    r.close()
    idx = np.arange(0, n, 128)
    m = mass.astype(np.float64)[:, None]
    np.savez_compressed(os.path.join(HERE, "c4_plummer2d_2e20_subsample.npz"), idx=idx, acc=acc[idx],
                        net_force=(m * acc).sum(axis=0), gross_force=np.abs(m * acc).sum(), softening=eps,
                        seconds=dt, n=n)
    print(f"reference compute_gravity at N=2^20: {dt:.1f} s ({n*(n-1)/dt:.3e} interactions/s)")

if __name__ == "__main__":
    main()
