"""Host-side plan for running the aggregator on several B200s of one box.

The reference has no parallelism on this path; the partition here is new.  Views are split evenly across
ranks for the whole stack (DINOv2, frame blocks, every linear layer), so the residual stream never moves.
Only global attention exchanges data: each rank's K|V rows [T_local, 2048] bf16 are all-gathered in rank
order into [world * T_local, 2048] and attended by the local queries.  Softmax is permutation-invariant over
keys and positions carry no view index, so rank-major order is exact.
"""
from __future__ import annotations

from typing import Optional, Sequence, Tuple

import torch


def view_shard(n_views: int, rank: int, world: int) -> Tuple[int, int]:
    """(first_view, count) of `rank`.  Equal shares only: attention FLOPs are balanced when views are, and the
    K|V all-gather wants equal counts.  Pad the view list by whole views upstream if it does not divide."""
    if world < 1 or not 0 <= rank < world:
        raise ValueError(f"bad rank/world {rank}/{world}")
    if n_views % world != 0:
        raise ValueError(f"{n_views} views do not divide over {world} ranks; pad by whole views")
    per = n_views // world
    return rank * per, per


def shard_inputs(images: torch.Tensor, priors, rank: int, world: int):
    """Slice [B, S, ...] inputs to this rank's views.  priors = (depth, rays, poses) or None."""
    first, cnt = view_shard(images.shape[1], rank, world)
    sl = slice(first, first + cnt)
    if priors is None:
        return images[:, sl], None, first
    return images[:, sl], tuple(None if p is None else p[:, sl] for p in priors), first


def kv_row_range(tokens_local: int, rank: int) -> Tuple[int, int]:
    """Rows of the gathered K|V buffer that belong to `rank` (what ncclAllGather's in-place layout implies)."""
    return rank * tokens_local, (rank + 1) * tokens_local


def broadcast_unique_id(uid: Optional[bytes], group=None, device: Optional[torch.device] = None) -> bytes:
    """Rank 0 passes the 128-byte NCCL id, the others None; every rank returns rank 0's bytes."""
    import torch.distributed as dist

    rank = dist.get_rank(group)
    t = torch.zeros(128, dtype=torch.uint8)
    if rank == 0:
        if uid is None or len(uid) != 128:
            raise ValueError("rank 0 must supply the 128-byte id")
        t = torch.tensor(list(uid), dtype=torch.uint8)
    if device is not None:
        t = t.to(device)
    src = dist.get_global_rank(group, 0) if group is not None else 0
    dist.broadcast(t, src=src, group=group)
    return bytes(t.cpu().tolist())


def gather_outputs(local: Sequence[torch.Tensor], group=None):
    """All-gather the per-rank [B, S_local, P, 2048] outputs along the view axis (needed only by callers that
    want every view on every rank, e.g. the camera head which attends across views, camera_head.py:70-92)."""
    import torch.distributed as dist

    world = dist.get_world_size(group)
    out = []
    for t in local:
        parts = [torch.empty_like(t) for _ in range(world)]
        dist.all_gather(parts, t.contiguous(), group=group)
        out.append(torch.cat(parts, dim=1))
    return out
