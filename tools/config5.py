"""Config 5 (BASELINE.json): 2^22-body two-galaxy collision, Velocity Verlet, target-partitioned over
the ranks of one box: strong-scaling time per step and energy drift from the fused diagnostics.
    python -m torch.distributed.run --nproc-per-node 8 --master-addr 127.0.0.1 tools/config5.py [n] [steps]
One JSON line on stdout from rank 0."""
import json, os, sys
ROOT = os.path.dirname(os.path.dirname(os.path.abspath(__file__))); sys.path.insert(0, ROOT)
import numpy as np
import __graft_entry__ as ge

def main():
    out = os.fdopen(os.dup(1), "w"); os.dup2(2, 1)
    n = int(sys.argv[1]) if len(sys.argv) > 1 else 1 << 22
    steps = int(sys.argv[2]) if len(sys.argv) > 2 else 10
    rank = int(os.environ.get("RANK", "0")); local = int(os.environ.get("LOCAL_RANK", "0")); world = int(os.environ.get("WORLD_SIZE", "1"))
    pkg = ge.load_package()
    pos, vel, mass = pkg.scenarios.two_galaxies(n, seed=1234, dim=3)
    eps = 3.844e8 / np.sqrt(n / 2)
    dt = 1e-3 * pkg.scenarios.crossing_time(n // 2)
    dist = None
    if world > 1:
        import torch, torch.distributed as dist
        torch.cuda.set_device(local)
        dist.init_process_group("cpu:gloo,cuda:nccl", device_id=torch.device("cuda", local))
        ph = pkg.sharded.create(dist, local)
    else:
        ph = pkg.Physics(local)
    ph.set_precision(pkg.PRECISION_FP32); ph.set_config(softening=eps, integrator=1)
    ph.upload(pos, vel, mass)
    ph.step(dt, 1, diagnostics=True); d0 = ph.step_diagnostics()
    ph.step(dt, 1); ph.synchronize()
    if dist is not None: dist.barrier()
    ph.timer_begin(); ph.step(dt, steps); ms = ph.timer_end()
    if dist is not None: ms = pkg.sharded.max_over_ranks(dist, ms)
    ph.step(dt, 1, diagnostics=True); d1 = ph.step_diagnostics()
    if rank == 0:
        out.write(json.dumps({"config": "5: two-galaxy collision", "n_bodies": n, "n_gpus": world, "steps": steps,
                              "ms_per_step": ms / steps, "interactions_per_s": float(n) * n * steps / (ms * 1e-3),
                              "energy0": d0.energy, "energy1": d1.energy, "steps_between": steps + 2,
                              "rel_energy_drift": (d1.energy - d0.energy) / abs(d0.energy),
                              "momentum_over_scale": float(np.abs(d1.momentum).max() / (mass.astype(np.float64).sum() * 1e3)),
                              "ok": bool(d0.ok and d1.ok)}) + "\n"); out.flush()
    ph.close()
    if dist is not None:
        dist.barrier(); dist.destroy_process_group()

if __name__ == "__main__":
    main()
