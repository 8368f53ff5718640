"""One-process-per-GPU plumbing for target-partitioned runs (torch.distributed is used for
rendezvous only: the id broadcast, result assembly and max-over-ranks timing; the per-step
exchange of packed positions is NCCL inside libnbody_b200.so).

Partitioning follows SURVEY.md §8(e): rank g owns a contiguous, tile-aligned slice of the
targets and every rank holds the full FP32 source stream.
"""
import numpy as np

from . import physics


def broadcast_unique_id(dist, make_id=physics.comm_unique_id):
    """Rank 0 draws the 128-byte NCCL id; everyone receives it."""
    box = [make_id() if dist.get_rank() == 0 else None]
    dist.broadcast_object_list(box, src=0)
    return box[0]


def create(dist, device):
    """A Physics context sharded over dist's world (plain single-GPU context when world == 1)."""
    rank, world = dist.get_rank(), dist.get_world_size()
    if world == 1:
        return physics.Physics(device)
    uid = broadcast_unique_id(dist)
    return physics.Physics(device, rank=rank, world=world, unique_id=uid)


def gather_rows(dist, local, n):
    """Assemble per-rank row slices (each rank passes an (n, dim) array with only its own rows
    filled, as Physics.download returns) into the full array on every rank."""
    import torch
    rank, world = dist.get_rank(), dist.get_world_size()
    lo, hi = physics.partition(n, rank, world)
    rows = max(physics.partition(n, r, world)[1] - physics.partition(n, r, world)[0] for r in range(world))
    dim = local.shape[1]
    buf = torch.zeros((rows, dim), dtype=torch.float64)
    buf[: hi - lo] = torch.from_numpy(np.ascontiguousarray(local[lo:hi]))
    parts = [torch.zeros_like(buf) for _ in range(world)]
    dist.all_gather(parts, buf)
    out = np.zeros((n, dim))
    for r in range(world):
        rlo, rhi = physics.partition(n, r, world)
        out[rlo:rhi] = parts[r][: rhi - rlo].numpy()
    return out


def max_over_ranks(dist, value):
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(dist, value):
    import torch
    t = torch.tensor([float(value)], dtype=torch.float64)
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return float(t.item())
