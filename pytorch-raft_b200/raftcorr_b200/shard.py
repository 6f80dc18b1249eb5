"""Batch sharding of frame pairs across ranks: every pair is independent, so the
multi-GPU path is one process per GPU, a contiguous slice of the batch per rank and no data-path
collective; only results (or timings) are gathered."""
from __future__ import annotations

from typing import List, Tuple


def shard_bounds(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced [lo, hi) slice of n_items for `rank` (first n_items % world ranks get one more)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    base, extra = divmod(n_items, world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def all_bounds(n_items: int, world: int) -> List[Tuple[int, int]]:
    return [shard_bounds(n_items, r, world) for r in range(world)]


def gather_pairs(local, n_items: int, group=None):
    """Reassemble per-rank results (tensors whose dim 0 is this rank's slice) on every rank.
    Uses all_gather on padded slices so it works for ragged shards with any backend."""
    import torch
    import torch.distributed as dist

    world = dist.get_world_size(group)
    bounds = all_bounds(n_items, world)
    width = max(hi - lo for lo, hi in bounds)
    pad = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: local.shape[0]] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, bounds)], dim=0)
