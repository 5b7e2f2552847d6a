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


def balanced_batch(n_scenes: int, workers: int, max_batch: int) -> int:
    """Scenes per tcs_segment_batch call for `n_scenes` scenes of one GPU served by `workers` contexts: at most
    `max_batch`, and sized so that every worker gets the same number of batches (128 scenes on 3 contexts: 6 batches
    of 21-22 instead of 4 of 32, which would leave two contexts idle in the second round).  The C++ driver uses the
    same rule (include/tcs/batch_segmenter.h)."""
    if workers < 1 or max_batch < 1:
        raise ValueError("bad workers/max_batch")
    if n_scenes <= 0:
        return 1
    rounds = max(1, -(-n_scenes // (workers * max_batch)))
    return max(1, min(max_batch, -(-n_scenes // (workers * rounds))))


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


# ---- label-column shard of ONE scene ------------------------------------------------------------
# Every rank builds the same graph and solves the colour columns columns_for_rank(K, rank, world) (the linear systems
# of different labels are independent); the only exchange is one all-gather of (best potential, its colour) per vertex,
# 8 bytes x V per rank - over NCCL when the ranks sit on different GPUs of the box.  The merged label of a vertex is the
# colour with the largest potential, the lowest colour on ties (the first-maximum rule of assignColors,
# impl/random_walker.hpp:209-222).  Scaling is sub-linear by construction: every rank re-reads the whole matrix.

def columns_for_rank(n_colours: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous block [begin, begin + count) of colour columns for this rank; blocks differ by at most one column."""
    if world < 1 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, extra = divmod(n_colours, world)
    begin = rank * base + min(rank, extra)
    return begin, base + (1 if rank < extra else 0)


def merge_best(best_val, best_col):
    """best_val / best_col: (world, V) arrays of per-shard best potential and colour -> merged (V,) colours."""
    import numpy as np
    val = np.asarray(best_val, np.float32)
    col = np.asarray(best_col, np.uint32)
    # largest potential wins; among equal potentials the lowest non-zero colour (colour 0 = unlabelled in that shard)
    # A vertex is labelled iff SOME shard labelled it (colour > 0: one of its potentials is non-zero - the rule of the
    # single-GPU k_argmax, which does not look at the sign); shards that left it unlabelled do not compete.
    key = np.where(col > 0, col, np.uint32(0xFFFFFFFF))
    val = np.where(col > 0, val, -np.inf)
    order = np.lexsort((key, -val), axis=0)[0]          # per vertex: shard with max val, then min colour
    c = np.take_along_axis(col, order[None, :], axis=0)[0]
    return c.astype(np.uint32)


def all_gather_best(val, col, world: int):
    """Gather the per-shard (V,) arrays of all ranks into (world, V) arrays with torch.distributed."""
    import numpy as np
    if world == 1:
        return np.asarray(val)[None, :], np.asarray(col)[None, :]
    import torch
    import torch.distributed as dist
    dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
    tv = torch.from_numpy(np.ascontiguousarray(val, np.float32)).to(dev)
    tc = torch.from_numpy(np.ascontiguousarray(col, np.uint32).view(np.int32)).to(dev)
    gv = [torch.empty_like(tv) for _ in range(world)]
    gc = [torch.empty_like(tc) for _ in range(world)]
    dist.all_gather(gv, tv)
    dist.all_gather(gc, tc)
    return (torch.stack(gv).cpu().numpy(), torch.stack(gc).cpu().numpy().view(np.uint32))


def segment_cloud_column_sharded(ctx, resolution, seed_xyz, seed_label, rank: int, world: int, preset=0, mode=0, opts=None):
    """One scene, label columns split over the ranks.  `ctx` has the cloud uploaded.  Returns (vertex colours merged
    over all ranks, this rank's result dict)."""
    import numpy as np
    from . import capi
    n_colours = len(np.unique(np.asarray(seed_label)))
    begin, count = columns_for_rank(n_colours, rank, world)
    opts = opts or capi.default_opts()
    opts.column_begin, opts.column_count = begin, max(count, 0)
    if count == 0:       # more ranks than colours: this rank contributes nothing
        V = None
        r = None
    else:
        r = ctx.segment_cloud(resolution, seed_xyz, seed_label, preset=preset, mode=mode, opts=opts)
        V = r["graph"].V
    if world > 1:
        import torch
        import torch.distributed as dist
        dev = "cuda" if dist.get_backend() == "nccl" else "cpu"
        t = torch.tensor([V or 0], dtype=torch.int64, device=dev)
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
        V = int(t.item())
    if world > 1:
        import torch.distributed as dist
        if dist.get_backend() == "nccl":
            # device-resident exchange: the all-gather reads the graph's own (best potential, colour) buffers - no
            # host staging; 8 bytes x V per rank over NVLink, then one merge pass on the device
            import torch
            if r is None:
                dev = torch.device("cuda", ctx.device)
                tv, tc = torch.zeros(V, dtype=torch.float32, device=dev), torch.zeros(V, dtype=torch.int32, device=dev)
            else:
                tv, tc = r["graph"].best_device()
            return _merge_best_nccl(tv, tc, world), r
    val, col = (np.zeros(V, np.float32), np.zeros(V, np.uint32)) if r is None else r["graph"].best()
    gv, gc = all_gather_best(val, col, world)
    return merge_best(gv, gc), r


def _merge_best_nccl(tv, tc, world: int):
    """tv (V,) f32 / tc (V,) i32 CUDA tensors of this rank -> merged (V,) uint32 colours on the host"""
    import numpy as np
    import torch
    import torch.distributed as dist
    gv = torch.empty((world,) + tv.shape, dtype=tv.dtype, device="cuda")
    gc = torch.empty((world,) + tc.shape, dtype=tc.dtype, device="cuda")
    dist.all_gather_into_tensor(gv, tv)
    dist.all_gather_into_tensor(gc, tc)
    gv = torch.where(gc > 0, gv, torch.full_like(gv, float("-inf")))   # unlabelled in a shard: does not compete
    best = gv.max(dim=0).values
    big = torch.iinfo(torch.int32).max
    cand = torch.where((gv == best[None, :]) & (gc > 0), gc, torch.full_like(gc, big))
    c = cand.min(dim=0).values
    c = torch.where(c != big, c, torch.zeros_like(c))
    return c.cpu().numpy().view(np.uint32)
