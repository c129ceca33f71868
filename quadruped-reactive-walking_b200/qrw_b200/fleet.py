"""Instance sharding over the GPUs of one node.

The robots of a fleet are independent: rank r of W owns the contiguous instance range
``shard_bounds(B, r, W)`` and keeps their persistent solver state on its own GPU.  There is no collective on
the solve path; the one exchange is the gather of the first-step contact forces (12 doubles per robot) at
the end of a tick -- ``gather_rows`` (torch.distributed all_gather: NCCL over NVLink on the GPU box, gloo in
the CPU tests).
"""
import numpy as np


def shard_bounds(total, rank, world):
    """[lo, hi) of rank's contiguous shard; the first ``total % world`` ranks take one extra instance."""
    if not (0 <= rank < world) or total < 0:
        raise ValueError("need 0 <= rank < world and total >= 0")
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def shard(array, rank, world):
    """Rows of an instance-major array owned by ``rank``."""
    lo, hi = shard_bounds(len(array), rank, world)
    return array[lo:hi]


def gather_rows(local, total, group=None):
    """All-gather instance-major rows ([n_local, ...] torch tensor) into [total, ...] on every rank.
    Shards may be ragged (total % world != 0): they are padded to the largest shard for the collective."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    if world == 1:
        return local
    biggest = shard_bounds(total, 0, world)[1]
    pad = torch.zeros((biggest,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    rows = []
    for r, p in enumerate(parts):
        lo, hi = shard_bounds(total, r, world)
        rows.append(p[: hi - lo])
    return torch.cat(rows, 0)


def first_step_forces(result):
    """[B, 24*n] MPC result (column-major 24 x n per instance) -> [B, 12] contact forces of step 0."""
    return result[:, 12:24]


__all__ = ["shard_bounds", "shard", "gather_rows", "first_step_forces", "np"]
