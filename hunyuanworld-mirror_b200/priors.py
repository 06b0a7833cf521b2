"""Prior preparation on the GPU: the per-pixel part of `WorldMirror.extract_priors`
(src/models/models/worldmirror.py:218-251).

`normalize_depth` mirrors `src/models/utils/priors.py:107-167` (same name, arguments and result) but normalises all
views with one kernel launch and no host synchronisation (the reference loops over views and calls `.item()` per view).
Pose normalisation and the intrinsics ratios touch S x 12 numbers and stay with the caller.  No CPU path."""
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
