"""Target-partitioned runs on >= 2 GPUs of one box (skipped on a single-GPU machine): the
sharded context (NCCL all-gather of packed positions overlapped with the own-sources pass,
rank-ordered FP64 diagnostics) must reproduce the single-GPU result."""
import os
import socket
import sys

import numpy as np
import pytest

from conftest import ROOT

pytestmark = pytest.mark.gpu


def _n_gpus():
    try:
        import torch
        return torch.cuda.device_count()
    except Exception:
        return 0


def _free_port():
    s = socket.socket()
    s.bind(("127.0.0.1", 0))
    port = s.getsockname()[1]
    s.close()
    return port


def _worker(rank, world, port, n, integ, out_dir):
    os.environ.update(MASTER_ADDR="127.0.0.1", MASTER_PORT=str(port))  # NCCL/gloo bootstrap address only
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    pkg = ge.load_package()
    torch.cuda.set_device(rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank),
                            init_method="file://" + os.path.join(out_dir, "rendezvous"))  # no TCP store port to collide on
    try:
        pos, vel, mass = pkg.scenarios.plummer(n, seed=17, dim=3)
        eps = 3.844e8 / np.sqrt(n)
        dt = 1e-3 * pkg.scenarios.crossing_time(n)
        ph = pkg.sharded.create(dist, rank)
        ph.set_precision(pkg.PRECISION_FP32)
        ph.set_config(softening=eps, integrator=integ)
        ph.upload(pos, vel, mass)
        ph.zero_net_momentum()
        ph.step(dt, 3)
        ph.step(dt, 2, diagnostics=True)
        d_fused = ph.step_diagnostics()
        d_alone = ph.compute_diagnostics()
        out = [pkg.sharded.gather_rows(dist, a, n) for a in ph.download()]
        lo, hi = ph.local_range()
        assert (lo, hi) == pkg.partition(n, rank, world)
        ph.close()
        if rank == 0:
            np.savez(os.path.join(out_dir, "sharded.npz"), pos=out[0], vel=out[1], acc=out[2],
                     fused=d_fused.as_reference_array(), alone=d_alone.as_reference_array(),
                     com=d_alone.com, mom=d_alone.momentum)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("integ", [1, 0])
@pytest.mark.parametrize("world,n", [(2, 20000), (2, 700), (8, 9000), (3, 5000)])
def test_sharded_matches_single_gpu(pkg, tmp_path, world, n, integ):
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    mp.spawn(_worker, args=(world, _free_port(), n, integ, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "sharded.npz")
    pos, vel, mass = pkg.scenarios.plummer(n, seed=17, dim=3)
    eps = 3.844e8 / np.sqrt(n)
    dt = 1e-3 * pkg.scenarios.crossing_time(n)
    ph = pkg.Physics(0)
    ph.set_precision(pkg.PRECISION_FP32)
    ph.set_config(softening=eps, integrator=integ)
    ph.upload(pos, vel, mass)
    ph.zero_net_momentum()
    ph.step(dt, 3)
    ph.step(dt, 2, diagnostics=True)
    d_fused = ph.step_diagnostics().as_reference_array()
    d_alone = ph.compute_diagnostics().as_reference_array()
    p, v, a = ph.download()
    ph.close()
    # Every rank orders its own slice of the stream along its own Hilbert curve, so the sharded run groups the FP32
    # pair terms differently from the single-GPU run (round 1 streamed in index order on every rank and agreed to
    # 1e-11): two valid FP32 evaluations of the same sums, a few 1e-7 of the largest acceleration apart.
    assert np.abs(got["acc"] - a).max() <= 1e-6 * np.abs(a).max()
    assert np.abs(got["pos"] - p).max() <= 1e-11 * np.abs(p).max()
    assert np.abs(got["vel"] - v).max() <= 1e-9 * np.abs(v).max()
    assert np.allclose(got["fused"][:3], d_fused[:3], rtol=1e-7, atol=0)
    assert np.allclose(got["alone"][:3], d_alone[:3], rtol=1e-7, atol=0)


# ---- one process, several GPUs (nb_create_multi): no torchrun, no NCCL -----------------------------
@pytest.mark.parametrize("integ", [1, 0])
@pytest.mark.parametrize("n_dev,n", [(2, 20000), (2, 700), (8, 9000), (3, 5000)])
def test_single_process_multi_gpu_matches_single_gpu(pkg, n_dev, n, integ):
    """The multi-GPU context a single-threaded host (the reference's flecs systems) can hold: members exchange
    the packed stream through peer memory. Same partition as the sharded context, so the same check: it must
    reproduce the single-GPU run (different summation grouping of the sources only), diagnostics included."""
    if _n_gpus() < n_dev:
        pytest.skip(f"needs {n_dev} GPUs")
    pos, vel, mass = pkg.scenarios.plummer(n, seed=17, dim=3)
    eps = 3.844e8 / np.sqrt(n)
    dt = 1e-3 * pkg.scenarios.crossing_time(n)
    out = []
    for devices in (None, list(range(n_dev))):
        ph = pkg.Physics(0) if devices is None else pkg.Physics(devices=devices)
        ph.set_precision(pkg.PRECISION_FP32)
        ph.set_config(softening=eps, integrator=integ)
        ph.upload(pos, vel, mass)
        ph.zero_net_momentum()
        ph.step(dt, 3)
        for _ in range(2):                      # per-frame calls: the speculative-drift path
            ph.step(dt, 1, diagnostics=True)
        fused = ph.step_diagnostics().as_reference_array()
        alone = ph.compute_diagnostics().as_reference_array()
        state = ph.download()
        lo, hi = ph.local_range()
        assert (lo, hi) == (0, n)               # one process sees every row
        ph.close()
        out.append((state, fused, alone))
    (s1, f1, a1), (s2, f2, a2) = out
    scale = np.abs(s1[0]).max()
    assert np.abs(s1[0] - s2[0]).max() / scale <= 1e-11
    assert np.abs(s1[1] - s2[1]).max() / np.abs(s1[1]).max() <= 1e-9
    assert np.abs(s1[2] - s2[2]).max() / np.abs(s1[2]).max() <= 1e-6
    assert np.abs(f1 - f2)[:3].max() <= 1e-7 * abs(f1[2]) and np.abs(a1 - a2)[:3].max() <= 1e-7 * abs(a1[2])


# ---- ownership along one curve (csrc/partition.cuh): the interface still speaks upload rows -----------------
@pytest.mark.parametrize("n_dev,n,dim", [(2, 5000, 3), (2, 700, 2), (3, 1500, 3), (8, 9000, 3), (2, 3, 2)])
def test_multi_gpu_upload_download_round_trip_is_exact(pkg, n_dev, n, dim):
    """Every GPU owns a stretch of one Hilbert curve of the whole world, so uploaded rows move between GPUs
    (csrc/partition.cuh). What the caller sees must not change: a download returns the uploaded rows, bit for bit,
    in the caller's order — also after a re-upload of a same-sized world, which re-uses the ownership."""
    if _n_gpus() < n_dev:
        pytest.skip(f"needs {n_dev} GPUs")
    rng = np.random.default_rng(n)
    ph = pkg.Physics(devices=list(range(n_dev)))
    ph.set_precision(pkg.PRECISION_FP32)
    for trial in range(3):
        pos = rng.standard_normal((n, dim)) * 1e8
        vel = rng.standard_normal((n, dim)) * 1e3
        acc = rng.standard_normal((n, dim))
        mass = rng.uniform(1e20, 1e24, n).astype(np.float32)
        pinned = (rng.uniform(size=n) < 0.1).astype(np.uint8)
        ph.upload_state(pos, vel, acc, mass, pinned)
        p, v, a, a0 = ph.download_state()
        assert np.array_equal(p, pos) and np.array_equal(v, vel) and np.array_equal(a, acc) and not a0.any()
    # pinned flags and masses followed their bodies: one Euler frame moves exactly the bodies that are not pinned
    ph.set_config(integrator=0, softening=1e7)
    ph.upload(pos, vel, mass, pinned)
    ph.step(1.0, 1)
    p1 = ph.download()[0]
    moved = np.any(p1 != pos, axis=1)
    assert np.array_equal(moved, pinned == 0)
    ph.close()


@pytest.mark.parametrize("integ", [1, 0])
def test_multi_gpu_checkpoint_resume_is_bit_exact(pkg, integ):
    """The stream order of a multi-GPU context names the caller's rows and fixes which GPU owns which body:
    state + stream order resume a run bit for bit in a fresh context (test_checkpoint_resume_is_bit_exact on 2 GPUs)."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    n = 6000
    pos, vel, mass = pkg.scenarios.plummer(n, seed=31, dim=3)
    eps = 3.844e8 / np.sqrt(n)
    dt = 1e-3 * pkg.scenarios.crossing_time(n)

    def fresh():
        ph = pkg.Physics(devices=[0, 1])
        ph.set_precision(pkg.PRECISION_FP32)
        ph.set_config(softening=eps, integrator=integ)
        return ph

    a = fresh()
    a.upload(pos, vel, mass)
    a.step(dt, 6)
    full = a.download_state()
    a.close()
    b = fresh()
    b.upload(pos, vel, mass)
    b.step(dt, 3)
    p, v, acc, acc_prev = b.download_state()
    order, age = b.stream_order()
    b.close()
    assert np.array_equal(np.sort(order), np.arange(n))  # a permutation of the caller's rows
    c = fresh()
    c.upload_state(p, v, acc if integ == 1 else None, mass)
    c.set_stream_order(order, age)
    c.step(dt, 3)
    resumed = c.download_state()
    c.close()
    for k in range(3):
        assert np.array_equal(resumed[k], full[k]), k


def _worker_round_trip(rank, world, n, out_dir):
    sys.path.insert(0, ROOT)
    import torch
    import torch.distributed as dist
    import __graft_entry__ as ge
    pkg = ge.load_package()
    torch.cuda.set_device(rank)
    dist.init_process_group("cpu:gloo,cuda:nccl", rank=rank, world_size=world, device_id=torch.device("cuda", rank),
                            init_method="file://" + os.path.join(out_dir, "rendezvous"))
    try:
        pos, vel, mass = pkg.scenarios.plummer(n, seed=5, dim=3)
        eps = 3.844e8 / np.sqrt(n)
        dt = 1e-3 * pkg.scenarios.crossing_time(n)
        ph = pkg.sharded.create(dist, rank)
        ph.set_precision(pkg.PRECISION_FP32)
        ph.set_config(softening=eps)
        acc0 = np.random.default_rng(1).standard_normal(pos.shape)
        ph.upload_state(pos, vel, acc0, mass)
        back = [pkg.sharded.gather_rows(dist, a, n) for a in ph.download_state()[:3]]
        exact = all(np.array_equal(x, y) for x, y in zip(back, (pos, vel, acc0)))
        # uninterrupted run / checkpoint after 3 frames / resume in a fresh context
        ph.upload(pos, vel, mass)
        ph.step(dt, 3)
        mid = [pkg.sharded.gather_rows(dist, a, n) for a in ph.download_state()]
        order, age = ph.stream_order()          # this rank's rows [lo,hi) only
        ph.step(dt, 3)
        full = [pkg.sharded.gather_rows(dist, a, n) for a in ph.download_state()]
        ph.close()
        ph = pkg.sharded.create(dist, rank)
        ph.set_precision(pkg.PRECISION_FP32)
        ph.set_config(softening=eps)
        ph.upload_state(mid[0], mid[1], mid[2], mass)
        ph.set_stream_order(order, age)
        ph.step(dt, 3)
        resumed = [pkg.sharded.gather_rows(dist, a, n) for a in ph.download_state()]
        ph.close()
        same = all(np.array_equal(x, y) for x, y in zip(resumed[:3], full[:3]))
        if rank == 0:
            np.savez(os.path.join(out_dir, "round_trip.npz"), exact=exact, same=same)
    finally:
        dist.destroy_process_group()


@pytest.mark.parametrize("world,n", [(2, 6000), (3, 1700)])
def test_sharded_round_trip_and_resume(pkg, tmp_path, monkeypatch, world, n):
    """One process per GPU: downloads return the uploaded rows exactly although the bodies live on other ranks,
    and state + per-rank stream order resume a run bit for bit."""
    if _n_gpus() < world:
        pytest.skip(f"needs {world} GPUs")
    import torch.multiprocessing as mp
    monkeypatch.setenv("NB_RESORT_EVERY", "4")   # the ranks regroup along the curve inside the 6-frame run
    mp.spawn(_worker_round_trip, args=(world, n, str(tmp_path)), nprocs=world, join=True)
    got = np.load(tmp_path / "round_trip.npz")
    assert bool(got["exact"]) and bool(got["same"])


@pytest.mark.parametrize("integ", [1, 0])
def test_multi_gpu_regroups_along_the_curve_during_a_run(pkg, monkeypatch, integ):
    """Long runs: at every re-sort of the stream (here every 3 generations instead of 64) the GPUs publish what
    they hold, recompute the curve from the current positions and regroup. The run still tracks the single-GPU
    run, downloads still return the caller's rows, and a checkpoint taken between two regroupings resumes bit for
    bit (ties inside one curve cell are broken by the caller's row, not by who held the body)."""
    if _n_gpus() < 2:
        pytest.skip("needs 2 GPUs")
    monkeypatch.setenv("NB_RESORT_EVERY", "3")
    n = 7000
    pos, vel, mass = pkg.scenarios.plummer(n, seed=23, dim=3)
    pinned = np.zeros(n, np.uint8)
    pinned[::97] = 1
    eps = 3.844e8 / np.sqrt(n)
    dt = 2e-3 * pkg.scenarios.crossing_time(n)

    def fresh(devices):
        ph = pkg.Physics(0) if devices is None else pkg.Physics(devices=devices)
        ph.set_precision(pkg.PRECISION_FP32)
        ph.set_config(softening=eps, integrator=integ)
        return ph

    one = fresh(None)
    one.upload(pos, vel, mass, pinned)
    one.step(dt, 11)
    ref = one.download_state()
    one.close()
    a = fresh([0, 1])
    a.upload(pos, vel, mass, pinned)
    for _ in range(11):
        a.step(dt, 1)                      # per-frame calls: speculative drifts in between
    full = a.download_state()
    info = a.partition_info()
    a.close()
    assert info["curve"] and info["partitions"] >= 3
    assert np.abs(full[0] - ref[0]).max() <= 1e-10 * np.abs(ref[0]).max()
    assert np.abs(full[1] - ref[1]).max() <= 1e-8 * np.abs(ref[1]).max()
    assert np.abs(full[2] - ref[2]).max() <= 1e-6 * np.abs(ref[2]).max()
    assert np.array_equal(full[0][pinned == 1], pos[pinned == 1])      # pins followed their bodies
    # checkpoint after 5 frames (between the regroupings at generations 3 and 6 ...), resume, 6 more frames
    b = fresh([0, 1])
    b.upload(pos, vel, mass, pinned)
    for _ in range(5):
        b.step(dt, 1)
    p, v, acc, _ = b.download_state()
    order, age = b.stream_order()
    b.close()
    c = fresh([0, 1])
    c.upload_state(p, v, acc if integ == 1 else None, mass, pinned)
    c.set_stream_order(order, age)
    for _ in range(6):
        c.step(dt, 1)
    resumed = c.download_state()
    c.close()
    for k in range(3):
        assert np.array_equal(resumed[k], full[k]), k
