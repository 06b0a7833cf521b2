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


def _role(key: str, shape: Tuple[int, ...]) -> Tuple[float, float]:
    """(mean, std) per tensor role."""
    leaf = key.rsplit(".", 1)[-1]
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
