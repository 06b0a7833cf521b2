"""CPU oracle for the Gaussian path of WorldMirror.forward.  TEST INFRASTRUCTURE ONLY
(same import rule as the other oracles: tests/, __graft_entry__.smoke() and bench.py's CPU legs).

Restates, from scratch, what `WorldMirror._gen_all_preds` (worldmirror.py:202-217) adds to the predictions when
`enable_gs` is set.  `GaussianSplatRenderer.render` (src/models/models/rasterization.py:166-246) rasterises with gsplat
and then DROPS the images; what survives is `gs_depth`, `gs_depth_conf` and `splats`, so this file covers
  gs_dpt_head      DPTHead with is_gsdpt (heads/dense_head.py:84-94, 232-245): depth / confidence + the feature map with
                   the 7x7 image branch added                                     (plain torch functional ops, fp32)
  gs_params        gs_renderer.gs_head: conv3x3 -> ReLU -> conv1x1 (rasterization.py:149-153)
  cameras          vector_to_camera_matrices + closed_form_inverse_se3 (utils/camera_utils.py:46-75, geometry.py:92-111)
  prepare_splats   rasterization.py:389-498 with position_from="gsdepth+predcamera", sh_degree 0, no offsets (the
                   configuration WorldMirror constructs, worldmirror.py:110-118); utils/act_gs.py, sh_utils.py:112
  prune_gs         rasterization.py:301-387 in NUMPY: np.unique + np.add.at, which adds in ascending index order exactly as
                   torch's CPU scatter_add_ does.

Parity status: PINNED against the reference itself.  tests/golden/make_golden_gs.py imports the real `DPTHead`,
`GaussianSplatRenderer` (gsplat's pure-python package on PYTHONPATH, its rasteriser stubbed: dead compute) from
/root/reference, runs config C1 with the synthetic weights and stores the results, plus the reference `prune_gs` on the
seeded clustered splats of `synthetic_splats` below, in tests/golden/gs_c1.npz; tests/test_gs_oracle.py checks this file
against that fixture (prune: bit-equal).
"""
from __future__ import annotations

from typing import Dict, Sequence, Tuple

import numpy as np
import torch
import torch.nn.functional as F

from . import heads_oracle as HO

Tensor = torch.Tensor
C0 = 0.28209479177387814  # utils/sh_utils.py:24
RAW_SPLIT = (4, 3, 1, 3, 1)  # quats | scales | opacities | residual SH (degree 0) | weights  (rasterization.py:131-138)


def gs_dpt_head(token_list: Sequence[Tensor], images: Tensor, sd: Dict[str, Tensor], pre: str = "gs_head.",
                patch_start: int = 7) -> Tuple[Tensor, Tensor, Tensor]:
    """DPTHead.forward with is_gsdpt=True, activation "exp+expp1".  images [1, S, 3, H, W].
    Returns (gs_feat [S, 128, H, W], gs_depth [1, S, H, W, 1], gs_depth_conf [1, S, H, W])."""
    H, W = images.shape[-2:]
    fused = HO.dpt_features(token_list, H, W, sd, pre, patch_start)
    o = F.conv2d(fused, sd[pre + "scratch.output_conv2.0.weight"], sd[pre + "scratch.output_conv2.0.bias"], padding=1)
    o = F.conv2d(F.relu(o), sd[pre + "scratch.output_conv2.2.weight"], sd[pre + "scratch.output_conv2.2.bias"])
    depth, conf = HO.activate(o, "exp+expp1")
    img_feat = F.relu(F.conv2d(images[0], sd[pre + "input_merger.0.weight"], sd[pre + "input_merger.0.bias"], padding=3))
    return fused + img_feat, depth[None], conf[None]


def gs_params(gs_feat: Tensor, sd: Dict[str, Tensor], pre: str = "gs_renderer.gs_head.") -> Tensor:
    """[S, 128, H, W] -> raw Gaussian parameters [S, H, W, 12]."""
    h = F.relu(F.conv2d(gs_feat, sd[pre + "0.weight"], None, padding=1))
    return F.conv2d(h, sd[pre + "2.weight"], sd[pre + "2.bias"]).permute(0, 2, 3, 1).contiguous()


def quat_to_rotmat(q: Tensor) -> Tensor:
    """utils/rotation.py:8-38, scalar-last quaternion."""
    i, j, k, r = q.unbind(-1)
    two_s = 2.0 / (q * q).sum(-1)
    o = torch.stack((1 - two_s * (j * j + k * k), two_s * (i * j - k * r), two_s * (i * k + j * r),
                     two_s * (i * j + k * r), 1 - two_s * (i * i + k * k), two_s * (j * k - i * r),
                     two_s * (i * k - j * r), two_s * (j * k + i * r), 1 - two_s * (i * i + j * j)), -1)
    return o.reshape(q.shape[:-1] + (3, 3))


def cameras(camera_params: Tensor, H: int, W: int) -> Tuple[Tensor, Tensor, Tensor]:
    """camera_params [S, 9] = [t, quat xyzw, fov_v, fov_u] -> (R_c2w [S,3,3], t_c2w [S,3], (fx, fy, cx, cy) [S,4])."""
    t, q = camera_params[:, :3], camera_params[:, 3:7]
    R = quat_to_rotmat(q)
    Rt = R.transpose(1, 2)
    t_inv = -(Rt @ t[:, :, None])[:, :, 0]
    fy = H * 0.5 / torch.tan(camera_params[:, 7] * 0.5)
    fx = W * 0.5 / torch.tan(camera_params[:, 8] * 0.5)
    k = torch.stack([fx, fy, torch.full_like(fx, W * 0.5), torch.full_like(fx, H * 0.5)], dim=-1)
    return Rt, t_inv, k


def prepare_splats(raw: Tensor, images: Tensor, gs_depth: Tensor, camera_params: Tensor) -> Dict[str, Tensor]:
    """raw [S, H, W, 12], images [S, 3, H, W], gs_depth [S, H, W], camera_params [S, 9] -> splats over N = S*H*W."""
    S, H, W, _ = raw.shape
    N = S * H * W
    quats, scales, opac, res_sh, weights = torch.split(raw.reshape(N, -1), list(RAW_SPLIT), dim=-1)
    out = {"quats": quats / (quats.norm(dim=-1, keepdim=True) + 1e-8),
           "scales": scales.exp().clamp_max(0.3),
           "opacities": opac[:, 0].sigmoid(),
           "weights": weights[:, 0].sigmoid()}
    rgb = images.permute(0, 2, 3, 1).reshape(N, 3)
    out["sh"] = (rgb - 0.5) / C0 + res_sh
    R, t, k = cameras(camera_params, H, W)
    v, u = torch.meshgrid(torch.arange(H, dtype=torch.float32), torch.arange(W, dtype=torch.float32), indexing="ij")
    z = gs_depth
    x = (u[None] - k[:, 2].view(S, 1, 1)) * z / k[:, 0].view(S, 1, 1)
    y = (v[None] - k[:, 3].view(S, 1, 1)) * z / k[:, 1].view(S, 1, 1)
    cam_pts = torch.stack([x, y, z], dim=-1)
    out["means"] = (torch.einsum("bhwi,bji->bhwj", cam_pts, R) + t[:, None, None, :]).reshape(N, 3)
    return out


def prune_gs(splats: Dict[str, np.ndarray], voxel_size: float = 0.002) -> Dict[str, np.ndarray]:
    """rasterization.py:301-387 for one scene, numpy fp32.  Input arrays over N splats (means [N,3], quats [N,4],
    scales [N,3], opacities [N], sh [N,3], weights [N]); output over the K occupied voxels in ascending flat index."""
    f32 = np.float32
    means = splats["means"].astype(f32)
    vox = np.floor(means / f32(voxel_size)).astype(np.int64)
    vox = vox - vox.min(axis=0)
    dims = vox.max(axis=0) + 1
    flat = vox[:, 0] * dims[1] * dims[2] + vox[:, 1] * dims[2] + vox[:, 2]
    _, inverse = np.unique(flat, return_inverse=True)
    K = int(inverse.max()) + 1
    w = splats["weights"].astype(f32)

    def wsum(x):  # scatter_add_(0, inverse, x * w), columns one by one
        x = x.astype(f32)
        x2 = x.reshape(x.shape[0], -1)
        acc = np.zeros((K, x2.shape[1]), dtype=f32)
        for d in range(x2.shape[1]):
            np.add.at(acc[:, d], inverse, (x2[:, d] * w).astype(f32))
        return acc

    tot = np.zeros(K, dtype=f32)
    np.add.at(tot, inverse, w)
    tot = np.maximum(tot, f32(1e-8))
    out = {"means": wsum(means) / tot[:, None], "sh": wsum(splats["sh"]) / tot[:, None],
           "scales": wsum(splats["scales"]) / tot[:, None]}
    op = np.zeros(K, dtype=f32)
    np.add.at(op, inverse, (w * w).astype(f32))
    out["opacities"] = op / tot
    q = wsum(splats["quats"])
    qn = np.sqrt((q * q).sum(axis=1, dtype=f32), dtype=f32)
    out["quats"] = q / np.maximum(qn, f32(1e-8))[:, None]
    return out


FILTER_KEYS = ("means", "quats", "scales", "opacities", "sh", "weights")


def conf_filter_count(n: int, percent: float = 30.0, max_gaussians: int = 5000000) -> int:
    """K of apply_confidence_filter (rasterization.py:272-276)."""
    keep = int(np.ceil(n * (100.0 - percent) / 100.0)) if percent > 0 else n
    return max(1, min(max_gaussians, keep))


def conf_filter_indices(conf: np.ndarray, percent: float = 30.0, max_gaussians: int = 5000000) -> np.ndarray:
    """Which splats apply_confidence_filter keeps (rasterization.py:266-279), one scene: confidences <= 1e-5 become -inf,
    the K largest survive.  torch.topk(sorted=False) leaves both the order of its result and the choice among equal values
    unspecified; stated here deterministically: equal values in ascending index, result in ascending index."""
    c = conf.reshape(-1).astype(np.float32).copy()
    c[c <= np.float32(1e-5)] = -np.inf
    K = conf_filter_count(c.shape[0], percent, max_gaussians)
    nan = np.isnan(c)  # torch.topk ranks NaN above everything (confidences are 1 + exp(.): NaN only from broken inputs)
    order = np.lexsort((np.arange(c.shape[0]), -np.where(nan, np.float32(np.inf), c), ~nan))  # NaN first, then descending, then index
    return np.sort(order[:K])


def apply_confidence_filter(splats: Dict[str, np.ndarray], conf: np.ndarray, percent: float = 30.0,
                            max_gaussians: int = 5000000) -> Dict[str, np.ndarray]:
    """rasterization.py:248-299 for one scene: the keys of `mask_keys` gathered at the kept indices, the rest untouched."""
    idx = conf_filter_indices(conf, percent, max_gaussians)
    return {k: (v[idx] if k in FILTER_KEYS else v) for k, v in splats.items()}


def synthetic_splats(n: int = 20000, seed: int = 5) -> Dict[str, np.ndarray]:
    """Seeded splats whose means cluster (many voxel collisions at voxel sizes 0.002 .. 0.05), negative coordinates
    included, a few exactly repeated points, a few zero weights."""
    g = np.random.default_rng(seed)
    centres = g.uniform(-0.5, 0.5, size=(n // 8, 3))
    means = centres[g.integers(0, centres.shape[0], size=n)] + g.normal(0.0, 0.004, size=(n, 3))
    means[::97] = means[0]
    q = g.normal(size=(n, 4))
    w = 1.0 / (1.0 + np.exp(-g.normal(-2.0, 1.0, size=n)))
    w[::211] = 0.0
    f = np.float32
    return {"means": means.astype(f), "quats": (q / np.linalg.norm(q, axis=1, keepdims=True)).astype(f),
            "scales": np.minimum(np.exp(g.normal(-5.0, 1.0, size=(n, 3))), 0.3).astype(f),
            "opacities": (1.0 / (1.0 + np.exp(-g.normal(size=n)))).astype(f),
            "sh": g.normal(size=(n, 3)).astype(f), "weights": w.astype(f)}


PRUNE_VOXELS = (0.002, 0.01, 0.05)


def gs_branch(token_list: Sequence[Tensor], images: Tensor, camera_params: Tensor, sd: Dict[str, Tensor],
              voxel_size: float = 0.002) -> Dict[str, object]:
    """What `_gen_all_preds` adds with enable_gs (worldmirror.py:202-217): gs_depth, gs_depth_conf and the pruned splats of
    one scene.  images [1, S, 3, H, W], camera_params [1, S, 9]."""
    feat, depth, conf = gs_dpt_head(token_list, images, sd)
    sp = prepare_splats(gs_params(feat, sd), images[0], depth[0, ..., 0], camera_params[0])
    return {"gs_depth": depth, "gs_depth_conf": conf, "splats": prune_gs({k: v.numpy() for k, v in sp.items()}, voxel_size)}
