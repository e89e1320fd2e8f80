"""Stream sharding across ranks: contiguous channel ranges, no data-path collective
(SURVEY.md section 8e).  The only exchange is a sum of small host-side counters."""
from __future__ import annotations


def shard_range(n_channels: int, world: int, rank: int) -> tuple[int, int]:
    """Channels [lo, hi) owned by `rank`: ceil(n/world) per rank, the last ranks may own fewer."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError("bad rank/world")
    per = -(-n_channels // world)
    lo = min(n_channels, rank * per)
    return lo, min(n_channels, lo + per)


def owner_of(channel: int, n_channels: int, world: int) -> int:
    per = -(-n_channels // world)
    return channel // per


def gather_counters(local: dict, world: int, device=None) -> dict:
    """Sum integer counters over all ranks (torch.distributed all_reduce; identity for world 1)."""
    if world <= 1:
        return dict(local)
    import torch
    import torch.distributed as dist
    keys = sorted(local)
    t = torch.tensor([int(local[k]) for k in keys], dtype=torch.int64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.SUM)
    return {k: int(v) for k, v in zip(keys, t.tolist())}
