"""World-size-2 CPU tests (gloo) of the host-side plumbing of target-partitioned runs:
id broadcast, partition agreement across ranks, assembly of per-rank row slices and the
max-over-ranks timing reduction. The per-step NCCL exchange itself is covered by the -m gpu
tests / bench at N > 1."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, out_dir):
    os.environ["MASTER_ADDR"] = "127.0.0.1"
    os.environ["MASTER_PORT"] = str(port)
    sys.path.insert(0, ROOT)
    import torch.distributed as dist
    import __graft_entry__ as ge
    pkg = ge.load_package()
    dist.init_process_group("gloo", rank=rank, world_size=world,
                            init_method="file://" + os.path.join(out_dir, "rendezvous"))  # no TCP store port to collide on
    try:
        # rank 0 draws the id; everyone must end up with the same 128 bytes
        uid = pkg.sharded.broadcast_unique_id(dist, make_id=lambda: bytes(range(128)))
        assert uid == bytes(range(128))
        lo, hi = pkg.partition(n, rank, world)
        full = np.arange(n * 2, dtype=np.float64).reshape(n, 2)
        mine = np.zeros_like(full)
        mine[lo:hi] = full[lo:hi]          # what Physics.download returns on a sharded context
        got = pkg.sharded.gather_rows(dist, mine, n)
        assert np.array_equal(got, full)
        assert pkg.sharded.max_over_ranks(dist, 10.0 + rank) == 10.0 + world - 1
        assert pkg.sharded.sum_over_ranks(dist, hi - lo) == n
        np.save(os.path.join(out_dir, f"ok{rank}.npy"), np.array([lo, hi]))
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("n", [3, 1000, 4096 + 5])
def test_two_rank_gloo_plumbing(pkg, tmp_path, n):
    import torch.multiprocessing as mp
    port = _free_port()
    mp.spawn(_worker, args=(2, port, n, str(tmp_path)), nprocs=2, join=True)
    spans = [np.load(tmp_path / f"ok{r}.npy") for r in range(2)]
    assert spans[0][0] == 0 and spans[0][1] == spans[1][0] and spans[1][1] == n


def test_snapshot_round_trip_is_bit_exact(pkg, tmp_path):
    """Scenario snapshots (Scenario.hpp:14-50 content, on disk): every float comes back bit for bit."""
    sc = pkg.scenarios
    pos, vel, mass = sc.uniform_disk(500, seed=4, circular=True)
    pos[3, 0] = np.nan
    vel[4, 1] = -0.0
    pinned = (np.arange(500) % 7 == 0).astype(np.uint8)
    acc = np.random.default_rng(1).standard_normal(pos.shape)
    cfg = {"g": 6.6743e-11, "softening": 4.0, "integrator": 1, "max_substep": 1e9}
    path = str(tmp_path / "snap.npz")
    sc.save_snapshot(path, pos, vel, mass, pinned, acc=acc, config=cfg, frame=123)
    s = sc.load_snapshot(path)
    assert s["pos"].tobytes() == pos.tobytes() and s["vel"].tobytes() == vel.tobytes()
    assert s["mass"].tobytes() == mass.tobytes() and np.array_equal(s["pinned"], pinned)
    assert s["acc"].tobytes() == acc.tobytes() and s["acc_prev"] is None
    assert s["config"] == cfg and s["frame"] == 123
    with pytest.raises(KeyError):
        sc.save_snapshot(path, pos, vel, mass, config={"bh_theta": 0.5})
