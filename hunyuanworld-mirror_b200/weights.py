"""Deterministic synthetic weights with the reference aggregator's state_dict schema.

The reference `VisualGeometryTransformer` (src/models/models/visual_transformer.py:20) has
1222 tensors / 916.2 M parameters.  Pretrained weights need the network and the reference
constructor does not exist on the GPU box, so parity runs on weights produced here: every
tensor is drawn from a CPU `torch.Generator` seeded by crc32(key) ^ seed, with a scale chosen
per tensor role so that activations stay O(1) through all 72 blocks and every code path
(QK-norm affine, LayerScale, biases, special tokens) carries signal.  Keys and shapes are
exactly the reference's, so the same dict loads with `strict=True` into the
reference module (tests/golden/make_golden.py does that) and into the B200 path.
"""
from __future__ import annotations

import zlib
from collections import OrderedDict
from typing import Dict, Tuple

import torch

EMBED = 1024
HEADS = 16
HEAD_DIM = 64
MLP = 4096
PATCH = 14
DINO_GRID = 37  # 518 / 14: native pos_embed grid
NUM_REG = 4


def _block_shapes(prefix: str, qk_norm: bool) -> "OrderedDict[str, Tuple[int, ...]]":
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    s[prefix + "norm1.weight"] = (EMBED,)
    s[prefix + "norm1.bias"] = (EMBED,)
    s[prefix + "attn.qkv.weight"] = (3 * EMBED, EMBED)
    s[prefix + "attn.qkv.bias"] = (3 * EMBED,)
    if qk_norm:
        s[prefix + "attn.q_norm.weight"] = (HEAD_DIM,)
        s[prefix + "attn.q_norm.bias"] = (HEAD_DIM,)
        s[prefix + "attn.k_norm.weight"] = (HEAD_DIM,)
        s[prefix + "attn.k_norm.bias"] = (HEAD_DIM,)
    s[prefix + "attn.proj.weight"] = (EMBED, EMBED)
    s[prefix + "attn.proj.bias"] = (EMBED,)
    s[prefix + "ls1.gamma"] = (EMBED,)
    s[prefix + "norm2.weight"] = (EMBED,)
    s[prefix + "norm2.bias"] = (EMBED,)
    s[prefix + "mlp.fc1.weight"] = (MLP, EMBED)
    s[prefix + "mlp.fc1.bias"] = (MLP,)
    s[prefix + "mlp.fc2.weight"] = (EMBED, MLP)
    s[prefix + "mlp.fc2.bias"] = (EMBED,)
    s[prefix + "ls2.gamma"] = (EMBED,)
    return s


def state_dict_schema(depth: int = 24, dino_depth: int = 24) -> "OrderedDict[str, Tuple[int, ...]]":
    """Key -> shape, in the reference's own registration order (probe of `state_dict()`)."""
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    s["cam_token"] = (1, 2, 1, EMBED)
    s["reg_token"] = (1, 2, NUM_REG, EMBED)
    s["patch_embed.cls_token"] = (1, 1, EMBED)
    s["patch_embed.pos_embed"] = (1, 1 + DINO_GRID * DINO_GRID, EMBED)
    s["patch_embed.register_tokens"] = (1, NUM_REG, EMBED)
    s["patch_embed.mask_token"] = (1, EMBED)
    s["patch_embed.patch_embed.proj.weight"] = (EMBED, 3, PATCH, PATCH)
    s["patch_embed.patch_embed.proj.bias"] = (EMBED,)
    for i in range(dino_depth):
        s.update(_block_shapes(f"patch_embed.blocks.{i}.", qk_norm=False))
    s["patch_embed.norm.weight"] = (EMBED,)
    s["patch_embed.norm.bias"] = (EMBED,)
    s["pose_embed.0.weight"] = (EMBED, 7)
    s["pose_embed.0.bias"] = (EMBED,)
    s["pose_embed.2.weight"] = (EMBED, EMBED)
    s["pose_embed.2.bias"] = (EMBED,)
    s["depth_embed.proj.2.fc1.weight"] = (MLP, PATCH * PATCH)
    s["depth_embed.proj.2.fc1.bias"] = (MLP,)
    s["depth_embed.proj.2.fc2.weight"] = (EMBED, MLP)
    s["depth_embed.proj.2.fc2.bias"] = (EMBED,)
    s["ray_embed.0.weight"] = (EMBED, 4)
    s["ray_embed.0.bias"] = (EMBED,)
    s["ray_embed.2.weight"] = (EMBED, EMBED)
    s["ray_embed.2.bias"] = (EMBED,)
    for i in range(depth):
        s.update(_block_shapes(f"frame_blocks.{i}.", qk_norm=True))
    for i in range(depth):
        s.update(_block_shapes(f"global_blocks.{i}.", qk_norm=True))
    return s


def _dpt_shapes(prefix: str, out_dim: int, dim_in: int = 2 * EMBED, features: int = 256) -> "OrderedDict[str, Tuple[int, ...]]":
    """DPTHead(dim_in=2048, output_dim=out_dim) state_dict (dense_head.py:36-106), registration order."""
    oc = (256, 512, 1024, 1024)
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    s[prefix + "norm.weight"] = (dim_in,)
    s[prefix + "norm.bias"] = (dim_in,)
    for i, c in enumerate(oc):
        s[prefix + f"projects.{i}.weight"] = (c, dim_in, 1, 1)
        s[prefix + f"projects.{i}.bias"] = (c,)
    s[prefix + "resize_layers.0.weight"] = (oc[0], oc[0], 4, 4)  # ConvTranspose2d: [in, out, k, k]
    s[prefix + "resize_layers.0.bias"] = (oc[0],)
    s[prefix + "resize_layers.1.weight"] = (oc[1], oc[1], 2, 2)
    s[prefix + "resize_layers.1.bias"] = (oc[1],)
    s[prefix + "resize_layers.3.weight"] = (oc[3], oc[3], 3, 3)
    s[prefix + "resize_layers.3.bias"] = (oc[3],)
    for i, c in enumerate(oc):
        s[prefix + f"scratch.layer{i + 1}_rn.weight"] = (features, c, 3, 3)
    for r in (1, 2, 3, 4):
        p = prefix + f"scratch.refinenet{r}."
        s[p + "out_conv.weight"] = (features, features, 1, 1)
        s[p + "out_conv.bias"] = (features,)
        for u in ((1, 2) if r != 4 else (2,)):
            for c in (1, 2):
                s[p + f"resConfUnit{u}.conv{c}.weight"] = (features, features, 3, 3)
                s[p + f"resConfUnit{u}.conv{c}.bias"] = (features,)
    s[prefix + "scratch.output_conv1.weight"] = (features // 2, features, 3, 3)
    s[prefix + "scratch.output_conv1.bias"] = (features // 2,)
    s[prefix + "scratch.output_conv2.0.weight"] = (32, features // 2, 3, 3)
    s[prefix + "scratch.output_conv2.0.bias"] = (32,)
    s[prefix + "scratch.output_conv2.2.weight"] = (out_dim, 32, 1, 1)
    s[prefix + "scratch.output_conv2.2.bias"] = (out_dim,)
    return s


def _camera_shapes(prefix: str, dim: int = 2 * EMBED, trunk_depth: int = 4) -> "OrderedDict[str, Tuple[int, ...]]":
    """CameraHead(dim_in=2048) state_dict (camera_head.py:16-56), registration order."""
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    s[prefix + "init_token"] = (1, 1, 9)
    for i in range(trunk_depth):
        p = prefix + f"refine_net.{i}."
        s[p + "norm1.weight"] = (dim,)
        s[p + "norm1.bias"] = (dim,)
        s[p + "attn.qkv.weight"] = (3 * dim, dim)
        s[p + "attn.qkv.bias"] = (3 * dim,)
        s[p + "attn.proj.weight"] = (dim, dim)
        s[p + "attn.proj.bias"] = (dim,)
        s[p + "ls1.gamma"] = (dim,)
        s[p + "norm2.weight"] = (dim,)
        s[p + "norm2.bias"] = (dim,)
        s[p + "mlp.fc1.weight"] = (4 * dim, dim)
        s[p + "mlp.fc1.bias"] = (4 * dim,)
        s[p + "mlp.fc2.weight"] = (dim, 4 * dim)
        s[p + "mlp.fc2.bias"] = (dim,)
        s[p + "ls2.gamma"] = (dim,)
    s[prefix + "token_norm.weight"] = (dim,)
    s[prefix + "token_norm.bias"] = (dim,)
    s[prefix + "out_norm.weight"] = (dim,)
    s[prefix + "out_norm.bias"] = (dim,)
    s[prefix + "param_embed.weight"] = (dim, 9)
    s[prefix + "param_embed.bias"] = (dim,)
    s[prefix + "adapt_norm_gen.1.weight"] = (3 * dim, dim)
    s[prefix + "adapt_norm_gen.1.bias"] = (3 * dim,)
    s[prefix + "param_predictor.fc1.weight"] = (dim // 2, dim)
    s[prefix + "param_predictor.fc1.bias"] = (dim // 2,)
    s[prefix + "param_predictor.fc2.weight"] = (9, dim // 2)
    s[prefix + "param_predictor.fc2.bias"] = (9,)
    return s


def heads_schema() -> "OrderedDict[str, Tuple[int, ...]]":
    """Key -> shape of WorldMirror's prediction heads under their attribute names (worldmirror.py:66-98):
    cam_head (CameraHead), pts_head / depth_head / norm_head (DPTHead with output_dim 4 / 2 / 4)."""
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    s.update(_camera_shapes("cam_head."))
    s.update(_dpt_shapes("pts_head.", 4))
    s.update(_dpt_shapes("depth_head.", 2))
    s.update(_dpt_shapes("norm_head.", 4))
    return s


def gs_schema(features: int = 256, raw_channels: int = 12) -> "OrderedDict[str, Tuple[int, ...]]":
    """Key -> shape of the Gaussian branch (worldmirror.py:101-118): gs_head = DPTHead(is_gsdpt=True, output_dim=2), whose
    extra module is input_merger = Conv2d(3, 128, 7, 1, 3) (dense_head.py:91-94), and gs_renderer.gs_head =
    Conv2d(128, 256, 3, bias=False) / ReLU / Conv2d(256, 12, 1) (rasterization.py:149-153; sh_degree 0, no offsets)."""
    s: "OrderedDict[str, Tuple[int, ...]]" = OrderedDict()
    s.update(_dpt_shapes("gs_head.", 2))
    s["gs_head.input_merger.0.weight"] = (features // 2, 3, 7, 7)
    s["gs_head.input_merger.0.bias"] = (features // 2,)
    s["gs_renderer.gs_head.0.weight"] = (features, features // 2, 3, 3)
    s["gs_renderer.gs_head.2.weight"] = (raw_channels, features, 1, 1)
    s["gs_renderer.gs_head.2.bias"] = (raw_channels,)
    return s


# constant_ initialisation of the parameter head's bias (rasterization.py:139-147): scales -7, opacities -2, weights -2
GS_BIAS_INIT = (0.0, 0.0, 0.0, 0.0, -7.0, -7.0, -7.0, -2.0, 0.0, 0.0, 0.0, -2.0)


def synthetic_heads_state_dict(seed: int = 0, gs: bool = False) -> Dict[str, torch.Tensor]:
    """Deterministic head weights under the reference keys (same recipe as `synthetic_state_dict`); gs: with the
    Gaussian branch (gs_head.*, gs_renderer.*)."""
    sd = OrderedDict((k, synthetic_tensor(k, shp, seed)) for k, shp in heads_schema().items())
    if gs:
        for k, shp in gs_schema().items():
            sd[k] = synthetic_tensor(k, shp, seed)
        sd["gs_renderer.gs_head.2.bias"] += torch.tensor(GS_BIAS_INIT)
    return sd


def _role(key: str, shape: Tuple[int, ...]) -> Tuple[float, float]:
    """(mean, std) per tensor role."""
    leaf = key.rsplit(".", 1)[-1]
    if key.endswith("init_token"):
        return 0.0, 0.1
    if "resize_layers.0." in key or "resize_layers.1." in key:
        # ConvTranspose2d with kernel == stride: every output pixel sees one tap, fan-in = in_channels (dim 0)
        return (0.0, 0.02) if leaf == "bias" else (0.0, 1.0 / float(shape[0]) ** 0.5)
    if "_head." in key and not key.startswith("cam_head.") and leaf == "weight" and len(shape) == 4:
        fan_in = shape[1] * shape[2] * shape[3]
        return 0.0, 1.0 / float(fan_in) ** 0.5
    if key.endswith("mask_token"):
        return 0.0, 0.0
    if key in ("cam_token", "reg_token"):
        return 0.0, 0.5
    if key.endswith("cls_token") or key.endswith("register_tokens"):
        return 0.0, 0.2
    if key.endswith("pos_embed"):
        return 0.0, 0.1
    if leaf == "gamma":
        # DINOv2 LayerScale is O(1) (init_values=1.0, visual_transformer.py:121); the
        # aggregator's is init 0.01 (visual_transformer.py:63).
        return (1.0, 0.1) if key.startswith("patch_embed.") else (0.01, 0.003)
    if "norm" in key and leaf == "weight":
        return 1.0, 0.1
    if "norm" in key and leaf == "bias":
        return 0.0, 0.05
    if leaf == "bias":
        return 0.0, 0.02
    if leaf == "weight":
        fan_in = 1
        for d in shape[1:]:
            fan_in *= d
        return 0.0, 0.64 / float(fan_in) ** 0.5
    raise KeyError(key)


def synthetic_tensor(key: str, shape: Tuple[int, ...], seed: int = 0) -> torch.Tensor:
    mean, std = _role(key, shape)
    if std == 0.0:
        return torch.full(shape, mean, dtype=torch.float32)
    g = torch.Generator(device="cpu")
    g.manual_seed((zlib.crc32(key.encode()) ^ (seed * 0x9E3779B1)) & 0x7FFFFFFF)
    t = torch.randn(shape, generator=g, dtype=torch.float32)
    return t.mul_(std).add_(mean)


def synthetic_state_dict(seed: int = 0, depth: int = 24, dino_depth: int = 24) -> Dict[str, torch.Tensor]:
    """fp32 CPU tensors under the reference's keys (depth < 24 gives a truncated stack whose
    keys are a prefix-consistent subset of the full model's: block i is identical)."""
    return OrderedDict(
        (k, synthetic_tensor(k, shp, seed)) for k, shp in state_dict_schema(depth, dino_depth).items()
    )


def synthetic_images(n_views: int, h: int, w: int, seed: int = 1) -> torch.Tensor:
    """[1, N, 3, H, W] uniform [0,1)."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    return torch.rand(1, n_views, 3, h, w, generator=g, dtype=torch.float32)


def synthetic_priors(n_views: int, h: int, w: int, seed: int = 2):
    """(depth[1,N,H,W] in [0,1], rays[1,N,4], poses[1,N,7]) as `extract_priors`
    (worldmirror.py:218-251) would hand them to the aggregator: normalised depth, intrinsics
    as (fx/W, fy/H, cx/W, cy/H), pose as [t(3) in the unit cube, quat xyzw(4)]."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    depth = torch.rand(1, n_views, h, w, generator=g)
    rays = torch.tensor([0.8, 0.8 * w / h, 0.5, 0.5]).repeat(1, n_views, 1)
    rays = rays + 0.05 * torch.randn(1, n_views, 4, generator=g)
    t = torch.rand(1, n_views, 3, generator=g) * 2 - 1
    q = torch.randn(1, n_views, 4, generator=g)
    q = q / q.norm(dim=-1, keepdim=True)
    return depth, rays, torch.cat([t, q], dim=-1)


def synthetic_scene(n_views: int, h: int, w: int, seed: int = 2):
    """Raw geometric inputs of `WorldMirror.forward` as SURVEY.md section 8d specifies them for the benchmark:
    camera_pose [1,N,4,4] = random small-rotation SE(3) (axis-angle <= 0.5 rad, t ~ U(-1,1)^3), depthmap [1,N,H,W] ~
    U(0.5, 4.5), camera_intrinsics [1,N,3,3] = pinhole fx = fy = 0.8 W, cx = W/2, cy = H/2.  `extract_priors` turns them
    into the (depth, rays, poses) tensors the aggregator takes."""
    g = torch.Generator(device="cpu")
    g.manual_seed(seed)
    axis = torch.randn(n_views, 3, generator=g)
    axis = axis / axis.norm(dim=-1, keepdim=True)
    ang = torch.rand(n_views, 1, generator=g) * 0.5
    k = axis
    K = torch.zeros(n_views, 3, 3)
    K[:, 0, 1], K[:, 0, 2], K[:, 1, 0], K[:, 1, 2], K[:, 2, 0], K[:, 2, 1] = -k[:, 2], k[:, 1], k[:, 2], -k[:, 0], -k[:, 1], k[:, 0]
    R = torch.eye(3)[None] + torch.sin(ang)[:, :, None] * K + (1 - torch.cos(ang))[:, :, None] * (K @ K)
    pose = torch.eye(4).repeat(n_views, 1, 1)
    pose[:, :3, :3] = R
    pose[:, :3, 3] = torch.rand(n_views, 3, generator=g) * 2 - 1
    depth = torch.rand(1, n_views, h, w, generator=g) * 4.0 + 0.5
    intr = torch.tensor([[0.8 * w, 0.0, w / 2.0], [0.0, 0.8 * w, h / 2.0], [0.0, 0.0, 1.0]]).repeat(1, n_views, 1, 1)
    return pose[None].contiguous(), depth, intr.contiguous()
