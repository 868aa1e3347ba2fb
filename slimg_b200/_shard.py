"""Data-parallel sharding of a batch of independent images over ranks (one process per GPU).

The reference parallelises over files with rayon's par_iter (cli/src/commands/mod.rs:64-72, convert.rs:60,
crop.rs:112, extend.rs:114): images are independent, so there is NO data-path collective.  torch.distributed is
used only to agree on the timing (barrier, MAX over ranks) and to gather per-image statuses, exactly what bench.py
does under torchrun.  Works with backend "nccl" (GPU box) and "gloo" (CPU tests)."""
from __future__ import annotations

from typing import List, Sequence, Tuple


def shard_range(n_items: int, rank: int, world: int) -> Tuple[int, int]:
    """Contiguous, balanced shard [lo, hi) of n_items for `rank`: sizes differ by at most one, every item is
    owned by exactly one rank (config 3: 1024 images over 1/2/4/8 GPUs -> 1024/512/256/128 per rank)."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank {rank} / world {world}")
    base, extra = divmod(max(0, n_items), world)
    lo = rank * base + min(rank, extra)
    return lo, lo + base + (1 if rank < extra else 0)


def max_over_ranks(value: float, device=None) -> float:
    """MAX of a per-rank scalar (device time of the timed region) -- the job is as slow as its slowest rank."""
    import torch
    import torch.distributed as dist
    t = torch.tensor([float(value)], dtype=torch.float64, device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size() > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX)
    return float(t.item())


def gather_statuses(local: Sequence[int], n_items: int, device=None) -> List[int]:
    """Per-image status codes of the whole job on every rank (ErrorCollector analogue, mod.rs:113-145): a failed
    image fails alone, and every rank learns which ones did."""
    import torch
    import torch.distributed as dist
    if not (dist.is_available() and dist.is_initialized()) or dist.get_world_size() == 1:
        return list(local)
    world, rank = dist.get_world_size(), dist.get_rank()
    full = torch.zeros(n_items, dtype=torch.int32, device=device)
    lo, hi = shard_range(n_items, rank, world)
    if hi > lo:
        full[lo:hi] = torch.tensor(list(local), dtype=torch.int32, device=device)
    dist.all_reduce(full, op=dist.ReduceOp.SUM)      # shards are disjoint: the sum is the concatenation
    return [int(v) for v in full.tolist()]


def whole_job_throughput(units_per_rank: float, world: int, seconds_max: float) -> float:
    """BASELINE metric for N ranks: units all ranks processed / the slowest rank's time (weak scaling)."""
    return units_per_rank * world / seconds_max
