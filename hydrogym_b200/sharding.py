"""Env-batch sharding across GPUs: one process per GPU, no collective on the step path.

``VecEnv`` in the reference is a pure ``vmap`` (hydrogym/jax/env_core.py:609-613): environments never
interact, so rank r simply owns a contiguous slice of the global batch.  The only (optional)
collective is gathering the tiny per-step outputs (obs 64 reals + reward per env) for a learner.
"""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(global_envs: int, rank: int, world: int):
    """Contiguous slice [lo, hi) of the global env batch owned by ``rank`` (sizes differ by at most 1)."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    base, rem = divmod(int(global_envs), int(world))
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_env_outputs(local: torch.Tensor, global_envs: int, group=None) -> torch.Tensor:
    """All-gather per-env outputs (leading axis = local envs) into global env order on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return local
    sizes = [shard_range(global_envs, r, world) for r in range(world)]
    # all_gather wants equal shapes: pad every shard to the largest one, slice after
    width = max(hi - lo for lo, hi in sizes)
    padded = local.new_zeros((width,) + tuple(local.shape[1:]))
    padded[: local.shape[0]] = local
    parts = [torch.empty_like(padded) for _ in range(world)]
    dist.all_gather(parts, padded, group=group)
    return torch.cat([p[: hi - lo] for p, (lo, hi) in zip(parts, sizes)], dim=0)


def max_over_ranks(seconds: float, device=None, group=None) -> float:
    """Timing reduction used by the bench: the slowest rank defines the step time."""
    if not dist.is_initialized() or dist.get_world_size(group) == 1:
        return float(seconds)
    t = torch.tensor([seconds], dtype=torch.float64, device=device)
    dist.all_reduce(t, op=dist.ReduceOp.MAX, group=group)
    return float(t.item())
