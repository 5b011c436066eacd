"""View-batch sharding across GPUs: a batch of camera views is split into contiguous
ranges, one per rank; the scene is replicated; there is no data-path collective.  The only exchange is
a fixed-size counter record per rank (views, pixels, spans, overflow flags ...) summed after a batch."""
from __future__ import annotations


def view_range(nviews: int, rank: int, world: int) -> tuple[int, int]:
    """[lo, hi) of the views rank `rank` renders; ranges are contiguous, disjoint, cover 0..nviews and differ
    in size by at most one view (ragged batches are allowed, empty ranges too when nviews < world)."""
    if world <= 0 or not (0 <= rank < world):
        raise ValueError("bad rank/world")
    base, rem = divmod(max(nviews, 0), world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_counters(local: dict, group=None) -> dict:
    """Sum the per-rank counter records over the process group (torch.distributed, any backend)."""
    import torch
    import torch.distributed as dist

    names = sorted(local)
    t = torch.tensor([float(local[k]) for k in names], dtype=torch.float64)
    if dist.is_available() and dist.is_initialized():
        if dist.get_backend(group) == "nccl":
            t = t.cuda()
        dist.all_reduce(t, op=dist.ReduceOp.SUM, group=group)
    return {k: int(v) for k, v in zip(names, t.tolist())}
