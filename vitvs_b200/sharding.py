"""Multi-GPU partitioning of the hot path: frame pairs are independent units, so a batch is
cut into contiguous blocks of pairs, one block per rank (one process per GPU), with NO collective on the data
path.  The only exchange is the optional gather of the tiny per-pair results (6 doubles + 2 scalars)."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_bounds(n_units: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [lo, hi) block of rank; sizes differ by at most one, earlier ranks take the remainder."""
    if not (0 <= rank < world):
        raise ValueError("rank outside [0, world)")
    base, rem = divmod(n_units, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_rows(local: torch.Tensor, n_units: int, group=None) -> torch.Tensor:
    """All ranks contribute their block of rows ([hi-lo, ...]); every rank gets the [n_units, ...] whole,
    in unit order.  Works on gloo (CPU tensors) and nccl (CUDA tensors); ragged blocks are padded."""
    world = dist.get_world_size(group)
    rank = dist.get_rank(group)
    lo, hi = shard_bounds(n_units, rank, world)
    assert local.shape[0] == hi - lo, (local.shape, lo, hi)
    width = -(-n_units // world)
    pad = torch.zeros((width,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    pad[: hi - lo] = local
    parts = [torch.empty_like(pad) for _ in range(world)]
    dist.all_gather(parts, pad, group=group)
    rows = []
    for r, p in enumerate(parts):
        l, h = shard_bounds(n_units, r, world)
        rows.append(p[: h - l])
    return torch.cat(rows, dim=0)


def run_sharded(step_fn, goal, current, depth, group=None):
    """Run ``step_fn(goal_block, current_block, depth_block) -> twist[block, 6]`` on this rank's block of
    the batch and gather the twists of the whole batch (unit order preserved)."""
    n = goal.shape[0]
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    rank = dist.get_rank(group) if dist.is_initialized() else 0
    lo, hi = shard_bounds(n, rank, world)
    tw = step_fn(goal[lo:hi], current[lo:hi], depth[lo:hi])
    if world == 1:
        return tw
    return gather_rows(tw, n, group)
