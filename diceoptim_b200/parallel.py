"""Candidate sharding across GPUs: one process per GPU, model replicated, candidates
split into contiguous index ranges; the only exchange is one (value, global index) pair per rank for
the global which.max (first maximum, NaN skipped; on equal values the lower global index wins, which
with contiguous ranges is the lower rank).  Works over any torch.distributed backend: NCCL on the
GPUs, gloo in the CPU tests."""
from __future__ import annotations

import math


def shard_range(N: int, rank: int, world: int) -> tuple[int, int]:
    """[lo, hi) of rank's contiguous share of N candidates; sizes differ by at most one."""
    base, rem = divmod(N, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def merge_argmax(pairs):
    """pairs: iterable of (value, global_index), index < 0 meaning 'nothing comparable on that rank'.
    Returns the which.max winner (value, index) or (nan, -1)."""
    best_v, best_i = math.nan, -1
    for v, i in pairs:
        i = int(i)
        if i < 0 or v != v:
            continue
        if best_i < 0 or v > best_v or (v == best_v and i < best_i):
            best_v, best_i = float(v), i
    return best_v, best_i


def global_argmax(local_value: float, local_index: int, offset: int, device="cpu"):
    """All-gather each rank's local winner (index local to its shard, -1 if none) and merge.
    Every rank returns the same (value, global_index)."""
    import torch
    import torch.distributed as dist
    gi = float(offset + local_index) if local_index >= 0 else -1.0
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return merge_argmax([(local_value, gi)])
    mine = torch.tensor([local_value, gi], dtype=torch.float64, device=device)   # indices < 2^53 are exact
    out = [torch.empty_like(mine) for _ in range(dist.get_world_size())]
    dist.all_gather(out, mine)
    return merge_argmax([(float(t[0]), float(t[1])) for t in torch.stack(out).cpu()])


def merge_topk(entries, k):
    """entries: (value, global_index) from all ranks -> the k best by (value desc, index asc)."""
    ok = [(float(v), int(i)) for v, i in entries if int(i) >= 0 and v == v]
    ok.sort(key=lambda t: (-t[0], t[1]))
    return ok[:k]
