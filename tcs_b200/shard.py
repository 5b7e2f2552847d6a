"""Host-side sharding of the path's natural units over the GPUs of one box.

Independent scenes are the natural shard: scene i goes to rank i mod world (round robin), every rank runs the whole
pipeline for its scenes on its own GPU and there is NO data-path collective.  The only communication is the
bookkeeping of the benchmark: total vertices (sum) and the slowest rank's device time (max).
"""
from __future__ import annotations

from typing import List, Sequence, Tuple


def scenes_for_rank(n_scenes: int, rank: int, world: int) -> List[int]:
    """Round-robin assignment: scene i -> rank i mod world."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    return list(range(rank, n_scenes, world))


def combine(local_units: float, local_ms: float, world: int) -> Tuple[float, float]:
    """(sum of units over ranks, max of milliseconds over ranks) - uses torch.distributed when world > 1."""
    if world == 1:
        return float(local_units), float(local_ms)
    import torch
    import torch.distributed as dist
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    s = torch.tensor([float(local_units)], dtype=torch.float64, device=dev)
    m = torch.tensor([float(local_ms)], dtype=torch.float64, device=dev)
    dist.all_reduce(s, op=dist.ReduceOp.SUM)
    dist.all_reduce(m, op=dist.ReduceOp.MAX)
    return s.item(), m.item()


def throughput(units: Sequence[float], ms: Sequence[float]) -> float:
    """Whole-job units per second from per-rank totals: sum(units) / max(ms)."""
    return sum(units) / (max(ms) * 1e-3)
