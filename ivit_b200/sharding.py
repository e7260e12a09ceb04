"""Batch sharding across ranks: every image is independent in the frozen forward,
so a batch is split contiguously over the ranks, weights are replicated, and the only collective is
one gather of the float32 logits.  One process per GPU; backend NCCL on the B200 box, gloo in the
CPU tests."""
from __future__ import annotations

import torch
import torch.distributed as dist


def shard_range(total: int, rank: int, world: int) -> tuple[int, int]:
    """Contiguous [lo, hi) of `total` items owned by `rank` (sizes differ by at most one)."""
    if not 0 <= rank < world:
        raise ValueError("rank out of range")
    base, rem = divmod(total, world)
    lo = rank * base + min(rank, rem)
    return lo, lo + base + (1 if rank < rem else 0)


def gather_logits(local: torch.Tensor, total: int) -> torch.Tensor:
    """All ranks receive the [total, classes] logits in global image order.  Ragged shards (total
    not divisible by world) are padded to the largest shard for the collective and trimmed after."""
    if not dist.is_initialized() or dist.get_world_size() == 1:
        return local
    world, rank = dist.get_world_size(), dist.get_rank()
    sizes = [shard_range(total, r, world)[1] - shard_range(total, r, world)[0] for r in range(world)]
    if local.shape[0] != sizes[rank]:
        raise ValueError(f"rank {rank} holds {local.shape[0]} rows, expected {sizes[rank]}")
    mx = max(sizes)
    padded = local if local.shape[0] == mx else torch.cat(
        [local, local.new_zeros((mx - local.shape[0],) + tuple(local.shape[1:]))])
    out = torch.empty((world * mx,) + tuple(local.shape[1:]), dtype=local.dtype, device=local.device)
    dist.all_gather_into_tensor(out, padded.contiguous())
    if all(s == mx for s in sizes):
        return out
    return torch.cat([out[r * mx:r * mx + sizes[r]] for r in range(world)])
