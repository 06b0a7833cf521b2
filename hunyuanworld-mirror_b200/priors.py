"""Prior preparation on the GPU: the per-pixel part of `WorldMirror.extract_priors`
(src/models/models/worldmirror.py:218-251).

`normalize_depth` mirrors `src/models/utils/priors.py:107-167` (same name, arguments and result) but normalises all
views with one kernel launch and no host synchronisation (the reference loops over views and calls `.item()` per view).
`extract_priors` mirrors the whole method: pose normalisation (first-view-relative, scene scale), the quaternion
pose vector and the intrinsics ratios run as one more launch (csrc/priors.cu: pose_prior_kernel).  No CPU path."""
from __future__ import annotations

import torch

from . import capi


@torch.no_grad()
def normalize_depth(depth: torch.Tensor, eps: float = 1e-6, min_percentile: int = 1, max_percentile: int = 99) -> torch.Tensor:
    """depth [B, S, H, W] -> [B, S, H, W] in [0, 1] (percentile-based scaling per view)."""
    if depth.dim() != 4:
        raise ValueError(f"Expected depth of shape [B, S, H, W], got {tuple(depth.shape)}")
    if not depth.is_cuda:
        raise RuntimeError("normalize_depth needs a CUDA tensor (there is no CPU path)")
    B, S, H, W = depth.shape
    src = depth.to(torch.float32).contiguous()
    out = torch.empty_like(src)
    with torch.cuda.device(depth.device):
        capi.check(capi.lib().wm_normalize_depth(torch.cuda.current_stream().cuda_stream, src.data_ptr(), B * S, H * W, float(eps),
                                                 int(min_percentile), int(max_percentile), out.data_ptr()))
    return out


@torch.no_grad()
def _gather_views(t: torch.Tensor, group, world: int) -> torch.Tensor:
    """[B, S_local, ...] on every rank of the sequence-parallel group -> [B, world * S_local, ...] in rank order."""
    import torch.distributed as dist

    t = t.contiguous()
    parts = [torch.empty_like(t) for _ in range(world)]
    dist.all_gather(parts, t, group=group)
    return torch.cat(parts, dim=1)


@torch.no_grad()
def extract_priors(views: dict, group=None, world: int = 1, rank: int = 0):
    """`WorldMirror.extract_priors` (worldmirror.py:218-251): views -> (depths [B,S,H,W] | None, rays [B,S,4] | None,
    poses [B,S,7] | None), the tuple `VisualGeometryTransformer.forward` takes.  Two kernel launches per scene instead of
    the reference's per-view Python loops; nothing is read back to the host.
    world > 1 (views sharded over the ranks of `group`, every rank holding S_local views of ONE scene): the pose
    normalisation is relative to the scene's first view and scaled by the mean camera distance over ALL views
    (priors.py:4-105), so the S x 16 pose (and S x 9 intrinsics) numbers are all-gathered first and this rank's rows
    are returned -- per-rank normalisation would silently give every rank but the first different pose tokens."""
    h, w = views["img"].shape[-2:]
    depths = rays = poses = None
    lib = capi.lib()
    if "camera_pose" in views:
        cp = views["camera_pose"]
        if not cp.is_cuda:
            raise RuntimeError("extract_priors needs CUDA tensors (there is no CPU path)")
        ext = cp.to(torch.float32).contiguous()  # [B, S, 4, 4] (or [B, S, 3, 4])
        s_local = ext.shape[1]
        intr = None
        if "camera_intrinsics" in views:
            intr = views["camera_intrinsics"][:, :, :3, :3].to(torch.float32).contiguous()
        if world > 1:
            ext = _gather_views(ext, group, world)
            intr = None if intr is None else _gather_views(intr, group, world)
        B, S = ext.shape[:2]
        stride = ext.shape[-2] * ext.shape[-1]
        if intr is not None:
            rays = torch.empty(B, S, 4, dtype=torch.float32, device=ext.device)
        poses = torch.empty(B, S, 7, dtype=torch.float32, device=ext.device)
        with torch.cuda.device(ext.device):
            st = torch.cuda.current_stream().cuda_stream
            for b in range(B):
                capi.check(lib.wm_pose_prior(st, ext[b].data_ptr(), stride, None if intr is None else intr[b].data_ptr(), S, h, w, 0.1,
                                             poses[b].data_ptr(), None if rays is None else rays[b].data_ptr()))
        if world > 1:  # this rank's views
            sl = slice(rank * s_local, (rank + 1) * s_local)
            poses = poses[:, sl].contiguous()
            rays = None if rays is None else rays[:, sl].contiguous()
    elif "camera_intrinsics" in views:
        k = views["camera_intrinsics"][:, :, :3, :3].to(torch.float32)
        rays = torch.stack([k[:, :, 0, 0] / w, k[:, :, 1, 1] / h, k[:, :, 0, 2] / w, k[:, :, 1, 2] / h], dim=-1)
    if "depthmap" in views:
        depths = normalize_depth(views["depthmap"])
    return depths, rays, poses
