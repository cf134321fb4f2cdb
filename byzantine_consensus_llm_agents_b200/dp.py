"""Data parallelism over model replicas: one process per GPU, full weights each.

Two independent things shard with no data-path collective:
  * ROWS of one `batch_generate_json` call — every (agent, phase) generation depends only on its
    own prompt.  All ranks run the same game script (same seeds -> same prompts), each decodes the
    rows `shard_rows()` gives it, and the texts are exchanged host-side (`gather_rows`, an
    all_gather_object of a few KB).  Per-row results are batch-invariant and seeded by the row's
    GLOBAL index, so the gathered result equals the single-GPU one.
  * whole GAMES — one game per replica (bench.py --gpus N, weak scaling).
Only host objects cross ranks (gloo or NCCL object collectives); device collectives are not used.
"""
from __future__ import annotations

from typing import Callable, List, Optional, Sequence


def dist_info():
    """(rank, world) of the default process group, (0, 1) when torch.distributed is not in use."""
    try:
        import torch.distributed as dist
        if dist.is_available() and dist.is_initialized():
            return dist.get_rank(), dist.get_world_size()
    except Exception:
        pass
    return 0, 1


def shard_rows(n_rows: int, rank: int, world: int) -> List[int]:
    """Round-robin: replica r takes rows r, r+world, ... (balanced to within one row)."""
    return list(range(rank, n_rows, world))


def gather_rows(local: Sequence, n_rows: int, rank: int, world: int, group=None) -> list:
    """Inverse of shard_rows across the group: every rank gets all n_rows results in row order."""
    if world == 1:
        return list(local)
    import torch.distributed as dist
    parts: List[Optional[list]] = [None] * world
    dist.all_gather_object(parts, list(local), group=group)
    out = [None] * n_rows
    for r, part in enumerate(parts):
        idx = shard_rows(n_rows, r, world)
        if len(part) != len(idx):
            raise RuntimeError(f"replica {r} returned {len(part)} rows, expected {len(idx)}")
        for i, v in zip(idx, part):
            out[i] = v
    return out


def run_sharded(run_local: Callable[[List[int]], list], n_rows: int, group=None) -> list:
    """Decode rows data-parallel: `run_local(global_row_indices) -> results` on each replica."""
    rank, world = dist_info()
    mine = shard_rows(n_rows, rank, world)
    local = run_local(mine) if mine else []
    return gather_rows(local, n_rows, rank, world, group)


def reduce_stats(stats: dict, group=None) -> dict:
    """Sum counters over replicas; keys ending in '_s' (seconds) take the max."""
    rank, world = dist_info()
    if world == 1:
        return dict(stats)
    import torch.distributed as dist
    parts = [None] * world
    dist.all_gather_object(parts, stats, group=group)
    out = {}
    for k in stats:
        vals = [p[k] for p in parts]
        out[k] = max(vals) if k.endswith("_s") else sum(vals)
    return out
