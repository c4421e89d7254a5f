"""Multi-GPU sharding of the env range: one process per GPU, independent shards, no data-path
collective.  The only exchange is the sum of the 16-float episode-statistics vector (env_mujoco.py:129-135
success-rate bookkeeping) -- NCCL over NVLink on the GPU box, gloo in the CPU tests.

Global env numbering: rank r of a world of W owns env ids [r * n_per_rank, (r + 1) * n_per_rank); the id seeds the
per-env Philox streams, so a sharded run reproduces the single-process run env for env.
"""
from __future__ import annotations

import torch
import torch.distributed as dist

STAT_NAMES = ["episodes", "successes", "return_sum", "timeouts", "drops", "singular_exits", "nan_resets",
              "env_steps", "length_sum"]


def shard_range(rank: int, world: int, n_per_rank: int):
    """(env_offset, n_envs) of `rank`: a contiguous, disjoint slice of the global env numbering."""
    if not (0 <= rank < world) or n_per_rank <= 0:
        raise ValueError(f"bad shard: rank {rank} of {world}, {n_per_rank} envs per rank")
    return rank * n_per_rank, n_per_rank


def reduce_stats(stats: torch.Tensor, group=None) -> torch.Tensor:
    """Sum the per-rank statistics vector over all ranks (in place; no-op without a process group)."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(stats, op=dist.ReduceOp.SUM, group=group)
    return stats


def max_over_ranks(value: float, device=None, group=None) -> float:
    """Max of a per-rank scalar (device-timed milliseconds in bench.py)."""
    t = torch.tensor([float(value)], device=device)
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())


def gather_per_rank(vec: torch.Tensor, group=None):
    """The same small float vector from every rank as a list of lists (rank order); [vec] without a process group.  Used by
    bench.py for its per-rank table (step time, workload fractions): evidence, not a data-path exchange."""
    if dist.is_available() and dist.is_initialized() and dist.get_world_size(group) > 1:
        out = [torch.zeros_like(vec) for _ in range(dist.get_world_size(group))]
        dist.all_gather(out, vec, group=group)
        return [o.cpu().tolist() for o in out]
    return [vec.cpu().tolist()]
