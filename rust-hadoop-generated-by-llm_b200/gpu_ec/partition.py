"""Stripe -> GPU partitioning: stripes are independent, so rank r of G owns the
stripes i with i % G == r (round-robin) and no collective touches the data path. The only
cross-rank traffic is bench bookkeeping (max elapsed time, summed byte counts)."""
from __future__ import annotations

from typing import List


def stripes_for_rank(n_stripes: int, rank: int, world: int) -> List[int]:
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, n_stripes, world))


def count_for_rank(n_stripes: int, rank: int, world: int) -> int:
    return len(range(rank, n_stripes, world))


def aggregate(elapsed_ms: float, units: float, dist=None):
    """(max elapsed over ranks, summed units). `dist` is torch.distributed or None (single process)."""
    if dist is None or not dist.is_initialized() or dist.get_world_size() == 1:
        return elapsed_ms, units
    import torch
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    t = torch.tensor([elapsed_ms], dtype=torch.float64, device=dev)
    u = torch.tensor([units], dtype=torch.float64, device=dev)
    dist.all_reduce(t, op=dist.ReduceOp.MAX)
    dist.all_reduce(u, op=dist.ReduceOp.SUM)
    return float(t.item()), float(u.item())
