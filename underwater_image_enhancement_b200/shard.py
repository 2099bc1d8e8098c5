"""Image-index sharding and the one exchange step of the path.

Frames are independent: rank r of R owns the contiguous block [r*ceil(M/R), min(M, (r+1)*ceil(M/R))).
The only collective is an all-gather of fixed-size per-image records
    [S*9 float64 scores (total + 8 sub-scores per strategy) | best strategy id | image type id]
with the last shard padded by sentinel rows (best = -1) so that every rank contributes equal counts.
torch.distributed is plumbing here (NCCL on the GPU box, gloo in the CPU tests)."""
from __future__ import annotations

import torch
import torch.distributed as dist

SENTINEL = -1.0


def shard_range(total: int, world: int, rank: int) -> tuple[int, int]:
    per = (total + world - 1) // world
    lo = min(total, rank * per)
    return lo, min(total, lo + per)


def shard_capacity(total: int, world: int) -> int:
    return (total + world - 1) // world


def pack_records(scores: torch.Tensor, best: torch.Tensor, types: torch.Tensor, capacity: int) -> torch.Tensor:
    """[n,S,9] f64, [n] i32, [n] i32 -> [capacity, S*9+2] f64 (rows beyond n are sentinels)."""
    n = scores.shape[0]
    width = scores.shape[1] * scores.shape[2] + 2
    rec = torch.full((capacity, width), SENTINEL, dtype=torch.float64, device=scores.device)
    if n:
        rec[:n, :-2] = scores.reshape(n, -1)
        rec[:n, -2] = best.to(torch.float64)
        rec[:n, -1] = types.to(torch.float64)
    return rec


def gather_records(rec: torch.Tensor, group=None) -> torch.Tensor:
    """All-gather equal-sized record blocks; returns [world*capacity, width] on every rank."""
    world = dist.get_world_size(group) if dist.is_initialized() else 1
    if world == 1:
        return rec
    out = torch.empty((world * rec.shape[0], rec.shape[1]), dtype=rec.dtype, device=rec.device)
    dist.all_gather_into_tensor(out, rec.contiguous(), group=group)
    return out


def unpack_records(gathered: torch.Tensor, strategies: int):
    """Drop sentinel rows; returns (scores [M,S,9], best [M], types [M]) in global image order."""
    keep = gathered[:, -2] != SENTINEL
    g = gathered[keep]
    return (g[:, :-2].reshape(-1, strategies, 9), g[:, -2].to(torch.int32), g[:, -1].to(torch.int32))
