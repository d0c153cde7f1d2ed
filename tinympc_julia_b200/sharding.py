"""Multi-GPU plumbing of the batched solve path: instance sharding and the one collective of the path.

Problem instances are independent: rank r of G owns a contiguous block of instances, the problem
constants are replicated by ``setup``, and nothing is exchanged while the ADMM kernels run.  The only collective is
the gather of the per-instance (status, iterations) pairs for reporting — NCCL over NVLink on the GPU box, gloo in the
CPU tests.  Device time of a multi-rank measurement is the maximum over ranks.
"""
from __future__ import annotations

import numpy as np


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous block [lo, hi) of ``total`` instances owned by ``rank`` (sizes differ by at most one)."""
    if world < 1 or not (0 <= rank < world) or total < 0:
        raise ValueError("need 0 <= rank < world and total >= 0")
    base, extra = divmod(total, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def gather_status(status: np.ndarray, iters: np.ndarray, device=None):
    """All-gather the (status, iterations) int32 pairs of every rank's shard.

    Returns (status_all, iters_all) concatenated in rank order on every rank.  Shards may have different lengths
    (they are padded to the longest for the collective).  Without an initialised process group the inputs are
    returned unchanged.
    """
    import torch
    import torch.distributed as dist
    status = np.ascontiguousarray(status, dtype=np.int32)
    iters = np.ascontiguousarray(iters, dtype=np.int32)
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return status, iters
    world = dist.get_world_size()
    dev = device if device is not None else torch.device("cpu")
    n = torch.tensor([status.size], dtype=torch.int64, device=dev)
    sizes = [torch.zeros_like(n) for _ in range(world)]
    dist.all_gather(sizes, n)
    sizes = [int(s.item()) for s in sizes]
    nmax = max(sizes)
    mine = torch.full((2, nmax), -1, dtype=torch.int32, device=dev)
    mine[0, : status.size] = torch.from_numpy(status).to(dev)
    mine[1, : iters.size] = torch.from_numpy(iters).to(dev)
    out = torch.empty((world * 2, nmax), dtype=torch.int32, device=dev)     # concatenation along dim 0 (gloo and nccl)
    dist.all_gather_into_tensor(out, mine)
    out = out.cpu().numpy().reshape(world, 2, nmax)
    st = np.concatenate([out[r, 0, : sizes[r]] for r in range(world)])
    it = np.concatenate([out[r, 1, : sizes[r]] for r in range(world)])
    return st, it


def max_over_ranks(value: float, device=None) -> float:
    """Maximum of a per-rank scalar (device time of a step): the slowest rank defines a multi-GPU measurement."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return float(value)
    t = torch.tensor([float(value)], dtype=torch.float64, device=device if device is not None else torch.device("cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def sum_over_ranks(values, device=None):
    import torch
    import torch.distributed as dist
    vals = [float(v) for v in values]
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return vals
    t = torch.tensor(vals, dtype=torch.float64, device=device if device is not None else torch.device("cpu"))
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return [float(v) for v in t.tolist()]
