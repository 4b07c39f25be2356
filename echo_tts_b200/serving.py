"""Multi-GPU serving: independent prompts, data-parallel, one process per GPU.

The sampling path has no cross-prompt operation (every tensor carries a leading batch axis: reference
``model.py:215-268``, ``inference.py:473-558``), so prompts are dealt round-robin to ranks, weights are replicated and
there is NO collective on the data path; only the finished waveforms (host tensors of different lengths) are gathered
on rank 0.  ``torch.distributed`` is plumbing here: NCCL on the GPU box, gloo in the CPU tests.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence, TypeVar

import torch
import torch.distributed as dist

T = TypeVar("T")
R = TypeVar("R")


def shard_indices(n_items: int, rank: int, world: int) -> List[int]:
    """Round-robin deal: item i belongs to rank i % world (keeps per-rank loads within one item of each other)."""
    if not (0 <= rank < world):
        raise ValueError(f"rank {rank} outside world of {world}")
    return list(range(rank, n_items, world))


def run_sharded(items: Sequence[T], generate: Callable[[T, int], R], *, group=None, dst: int = 0) -> Optional[List[R]]:
    """Run ``generate(item, global_index)`` for this rank's share of ``items`` and gather the results, in the
    original order, on rank ``dst`` (other ranks get None).  Results must be picklable host objects (e.g. CPU
    waveforms): gathering is control-plane traffic, not part of the timed data path."""
    if not (dist.is_available() and dist.is_initialized()):
        return [generate(it, i) for i, it in enumerate(items)]
    rank, world = dist.get_rank(group), dist.get_world_size(group)
    mine = shard_indices(len(items), rank, world)
    local = [(i, generate(items[i], i)) for i in mine]
    gathered: Optional[list] = [None] * world if rank == dst else None
    dist.gather_object(local, gathered, dst=dst, group=group)
    if rank != dst:
        return None
    out: List[Optional[R]] = [None] * len(items)
    for part in gathered:
        for i, r in part:
            out[i] = r
    assert all(r is not None for r in out) or len(items) == 0
    return out  # type: ignore[return-value]


def max_over_ranks_ms(ms: float, device: Optional[torch.device] = None, group=None) -> float:
    """Device-timed milliseconds -> max over ranks (the number every multi-GPU figure is quoted on)."""
    if not (dist.is_available() and dist.is_initialized()):
        return ms
    t = torch.tensor([ms], dtype=torch.float64, device=device or "cpu")
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
