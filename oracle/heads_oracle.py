"""CPU oracle for the WorldMirror prediction heads.  TEST INFRASTRUCTURE ONLY.

A from-scratch fp32 restatement (plain torch functional ops, no nn.Module, no reference import) of
`CameraHead.forward` (src/models/heads/camera_head.py:58-104) and `DPTHead._forward_impl`
(src/models/heads/dense_head.py:164-260) as `WorldMirror._gen_all_preds` (worldmirror.py:154-196) calls them
with autocast disabled (fp32).  Same import rule as oracle/aggregator_oracle.py: only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s CPU-baseline legs may import this file.

Parity status: PINNED against the reference itself: `tests/golden/make_golden_heads.py` imports the real
`CameraHead` / `DPTHead` from /root/reference in the build container, loads the synthetic head weights
(hunyuanworld-mirror_b200/weights.py: `synthetic_heads_state_dict`, reference keys, strict=True), feeds them
the reference aggregator's own `token_list` for config C1 and stores the predictions in
tests/golden/c1_heads.npz; tests/test_heads_oracle.py checks this file against that fixture.

These functions are what turns the aggregator's `token_list` into the quantities BASELINE.json's north_star
states tolerances on (depth, pts3d, normals, camera rotation), so the GPU parity test can check those
tolerances end to end: B200 token_list -> this oracle's heads vs the reference's token_list -> reference heads.
"""
from __future__ import annotations

import math
from typing import Dict, List, Sequence, Tuple

import torch
import torch.nn.functional as F

from . import aggregator_oracle as A

Tensor = torch.Tensor
PATCH = 14


# --------------------------------------------------------------------------------------
# Camera head (camera_head.py)
# --------------------------------------------------------------------------------------
def _attention_plain(x: Tensor, sd: Dict[str, Tensor], pre: str, heads: int) -> Tensor:
    """Attention.forward without qk-norm / rope (attention.py:48-69); x: [B, L, C]."""
    B, L, C = x.shape
    d = C // heads
    qkv = A.linear(x, sd[pre + "qkv.weight"], sd[pre + "qkv.bias"]).reshape(B, L, 3, heads, d).permute(2, 0, 3, 1, 4)
    o = A.sdpa(qkv[0], qkv[1], qkv[2]).transpose(1, 2).reshape(B, L, C)
    return A.linear(o, sd[pre + "proj.weight"], sd[pre + "proj.bias"])


def _block_plain(x: Tensor, sd: Dict[str, Tensor], pre: str, heads: int) -> Tensor:
    """Block.forward eval branch (block.py:91-92), nn.LayerNorm default eps 1e-5, LayerScale init 0.01."""
    n1 = A.layer_norm(x, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], 1e-5)
    x = x + sd[pre + "ls1.gamma"] * _attention_plain(n1, sd, pre + "attn.", heads)
    n2 = A.layer_norm(x, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], 1e-5)
    return x + sd[pre + "ls2.gamma"] * A.mlp(n2, sd, pre + "mlp.")


def camera_activation(p: Tensor) -> Tensor:
    """apply_camera_parameter_activation (camera_head.py:106-126): trans linear, quat linear, fov relu."""
    return torch.cat([p[..., :7], F.relu(p[..., 7:])], dim=-1)


def camera_head(cam_tokens: Tensor, sd: Dict[str, Tensor], pre: str = "cam_head.", steps: int = 4,
                trunk_depth: int = 4, heads: int = 16) -> List[Tensor]:
    """CameraHead.forward (camera_head.py:58-104).  cam_tokens: [B, S, 2048] = token_list[-1][:, :, 0].
    Returns the activated [B, S, 9] prediction of every refinement step ([t(3), quat xyzw(4), fov(2)])."""
    tok = A.layer_norm(cam_tokens, sd[pre + "token_norm.weight"], sd[pre + "token_norm.bias"], 1e-5)
    B, S, C = tok.shape
    ones = torch.ones(C, dtype=tok.dtype)
    zeros = torch.zeros(C, dtype=tok.dtype)
    pred = None
    seq = []
    for _ in range(steps):
        inp = sd[pre + "init_token"].expand(B, S, -1) if pred is None else pred
        emb = A.linear(inp, sd[pre + "param_embed.weight"], sd[pre + "param_embed.bias"])
        mod = A.linear(emb * torch.sigmoid(emb), sd[pre + "adapt_norm_gen.1.weight"], sd[pre + "adapt_norm_gen.1.bias"])
        shift, scale, gate = mod.chunk(3, dim=-1)
        x = gate * (A.layer_norm(tok, ones, zeros, 1e-6) * (1 + scale) + shift) + tok  # :82-83, adaLN eps 1e-6 (:52)
        for i in range(trunk_depth):
            x = _block_plain(x, sd, f"{pre}refine_net.{i}.", heads)
        x = A.layer_norm(x, sd[pre + "out_norm.weight"], sd[pre + "out_norm.bias"], 1e-5)
        delta = A.mlp(x, sd, pre + "param_predictor.")
        pred = delta if pred is None else pred + delta
        seq.append(camera_activation(pred))
    return seq


def quat_angle(qa: Tensor, qb: Tensor) -> Tensor:
    """Rotation angle (rad) between the rotations of quaternions qa, qb [..., 4] (normalised here, sign-invariant).
    2 acos|<qa,qb>| loses everything below 1e-3 rad in fp32 (acos near 1), so the chord form
    4 asin(|qa - s qb| / 2), s = sign<qa,qb>, is evaluated in float64."""
    qa, qb = qa.double(), qb.double()
    qa = qa / qa.norm(dim=-1, keepdim=True)
    qb = qb / qb.norm(dim=-1, keepdim=True)
    s = torch.where((qa * qb).sum(-1, keepdim=True) < 0, -1.0, 1.0)
    return 4.0 * torch.asin(((qa - s * qb).norm(dim=-1) / 2.0).clamp(max=1.0))


# --------------------------------------------------------------------------------------
# DPT head (dense_head.py)
# --------------------------------------------------------------------------------------
def uv_pos_embed(w: int, h: int, channels: int, aspect: float, ratio: float = 0.1) -> Tensor:
    """_apply_pos_embed (dense_head.py:262-271) = create_uv_grid (utils/grid.py:46-89) ->
    position_grid_to_embed (utils/grid.py:4-42) * ratio.  Returns [channels, h, w] fp32.
    u spans +-span_x (w-1)/w, v spans +-span_y (h-1)/h; omega_k = 100^(-k / (C/4)) in float64;
    channel order [sin(u w), cos(u w), sin(v w), cos(v w)], each C/4 wide."""
    diag = math.sqrt(aspect * aspect + 1.0)
    sx, sy = aspect / diag, 1.0 / diag
    u = torch.linspace(-sx * (w - 1) / w, sx * (w - 1) / w, w, dtype=torch.float32)
    v = torch.linspace(-sy * (h - 1) / h, sy * (h - 1) / h, h, dtype=torch.float32)
    omega = torch.arange(channels // 4, dtype=torch.float64) / (channels / 4.0)
    omega = 1.0 / 100.0 ** omega
    au = u.double()[:, None] * omega[None]  # [w, C/4]
    av = v.double()[:, None] * omega[None]  # [h, C/4]
    eu = torch.cat([au.sin(), au.cos()], dim=1).float()  # [w, C/2]
    ev = torch.cat([av.sin(), av.cos()], dim=1).float()  # [h, C/2]
    emb = torch.cat([eu[None].expand(h, -1, -1), ev[:, None].expand(-1, w, -1)], dim=-1)  # [h, w, C]
    return (emb * ratio).permute(2, 0, 1)


def _rcu(x: Tensor, sd: Dict[str, Tensor], pre: str) -> Tensor:
    """ResidualConvUnit.forward (dense_head.py:436-457).  The activation is nn.ReLU(inplace=True), so the first
    `self.activation(x)` overwrites x: the skip connection adds relu(x), not x."""
    x = F.relu(x)
    out = F.conv2d(x, sd[pre + "conv1.weight"], sd[pre + "conv1.bias"], padding=1)
    out = F.conv2d(F.relu(out), sd[pre + "conv2.weight"], sd[pre + "conv2.bias"], padding=1)
    return out + x


def _fusion(sd: Dict[str, Tensor], pre: str, x: Tensor, skip, size) -> Tensor:
    """FeatureFusionBlock.forward (dense_head.py:506-536): [+ RCU1(skip)] -> RCU2 -> bilinear (align_corners)
    to `size` (x2 when None) -> 1x1 out_conv."""
    if skip is not None:
        x = x + _rcu(skip, sd, pre + "resConfUnit1.")
    x = _rcu(x, sd, pre + "resConfUnit2.")
    if size is None:
        size = (x.shape[-2] * 2, x.shape[-1] * 2)
    x = F.interpolate(x, size=tuple(size), mode="bilinear", align_corners=True)
    return F.conv2d(x, sd[pre + "out_conv.weight"], sd[pre + "out_conv.bias"])


def inv_log(x: Tensor) -> Tensor:
    """sign(y) (exp|y| - 1) (dense_head.py:346-356)."""
    return torch.sign(x) * torch.expm1(torch.abs(x))


def activate(out: Tensor, activation: str) -> Tuple[Tensor, Tensor]:
    """activate_head (dense_head.py:297-344) for the three activations WorldMirror uses (worldmirror.py:74-98):
    'inv_log+expp1' (pts), 'exp+expp1' (depth), 'norm+expp1' (normals).  out: [N, C, H, W] -> ([N,H,W,C-1], [N,H,W])."""
    act, conf_act = activation.split("+")
    assert conf_act == "expp1"
    feat = out.permute(0, 2, 3, 1)
    attr, conf = feat[..., :-1], feat[..., -1]
    if act == "inv_log":
        attr = inv_log(attr)
    elif act == "exp":
        attr = torch.exp(attr)
    elif act == "norm":
        attr = attr / attr.norm(dim=-1, keepdim=True)
    else:
        raise ValueError(f"Unknown attribute activation: {act}")
    return attr, 1 + conf.exp()


def dpt_features(token_list: Sequence[Tensor], H: int, W: int, sd: Dict[str, Tensor], pre: str,
                 patch_start: int = 7) -> Tensor:
    """DPTHead._forward_impl up to the fused full-resolution feature map (dense_head.py:186-232).
    token_list: 4 x [1, S, P, 2048].  Returns [S, 128, H, W] (after the second positional embedding)."""
    S = token_list[0].shape[1]
    ph, pw = H // PATCH, W // PATCH
    feats = []
    for i, tokens in enumerate(token_list):
        t = tokens[0, :, patch_start:]  # [S, g, 2048]
        t = A.layer_norm(t, sd[pre + "norm.weight"], sd[pre + "norm.bias"], 1e-5)
        f = t.permute(0, 2, 1).reshape(S, t.shape[-1], ph, pw)
        f = F.conv2d(f, sd[pre + f"projects.{i}.weight"], sd[pre + f"projects.{i}.bias"])
        f = f + uv_pos_embed(pw, ph, f.shape[1], W / H)[None]
        if i == 0:
            f = F.conv_transpose2d(f, sd[pre + "resize_layers.0.weight"], sd[pre + "resize_layers.0.bias"], stride=4)
        elif i == 1:
            f = F.conv_transpose2d(f, sd[pre + "resize_layers.1.weight"], sd[pre + "resize_layers.1.bias"], stride=2)
        elif i == 3:
            f = F.conv2d(f, sd[pre + "resize_layers.3.weight"], sd[pre + "resize_layers.3.bias"], stride=2, padding=1)
        feats.append(f)
    rn = [F.conv2d(f, sd[pre + f"scratch.layer{i + 1}_rn.weight"], None, padding=1) for i, f in enumerate(feats)]
    out = _fusion(sd, pre + "scratch.refinenet4.", rn[3], None, rn[2].shape[2:])
    out = _fusion(sd, pre + "scratch.refinenet3.", out, rn[2], rn[1].shape[2:])
    out = _fusion(sd, pre + "scratch.refinenet2.", out, rn[1], rn[0].shape[2:])
    out = _fusion(sd, pre + "scratch.refinenet1.", out, rn[0], None)
    out = F.conv2d(out, sd[pre + "scratch.output_conv1.weight"], sd[pre + "scratch.output_conv1.bias"], padding=1)
    out = F.interpolate(out, size=(ph * PATCH, pw * PATCH), mode="bilinear", align_corners=True)
    return out + uv_pos_embed(out.shape[-1], out.shape[-2], out.shape[1], W / H)[None]


def dpt_head(token_list: Sequence[Tensor], H: int, W: int, sd: Dict[str, Tensor], pre: str, activation: str,
             patch_start: int = 7) -> Tuple[Tensor, Tensor]:
    """DPTHead.forward, is_gsdpt=False (dense_head.py:108-260).  Returns (pred [1,S,H,W,C-1], conf [1,S,H,W])."""
    fused = dpt_features(token_list, H, W, sd, pre, patch_start)
    o = F.conv2d(fused, sd[pre + "scratch.output_conv2.0.weight"], sd[pre + "scratch.output_conv2.0.bias"], padding=1)
    o = F.conv2d(F.relu(o), sd[pre + "scratch.output_conv2.2.weight"], sd[pre + "scratch.output_conv2.2.bias"])
    pred, conf = activate(o, activation)
    return pred[None], conf[None]


HEAD_ACTIVATIONS = {"depth_head.": "exp+expp1", "pts_head.": "inv_log+expp1", "norm_head.": "norm+expp1"}


def all_heads(token_list: Sequence[Tensor], H: int, W: int, sd: Dict[str, Tensor]) -> Dict[str, Tensor]:
    """WorldMirror._gen_all_preds (worldmirror.py:154-196) without the Gaussian branch: camera_params, depth,
    pts3d, normals and their confidences."""
    preds = {"camera_params": camera_head(token_list[-1][:, :, 0], sd)[-1]}
    for pre, name in (("depth_head.", "depth"), ("pts_head.", "pts3d"), ("norm_head.", "normals")):
        p, c = dpt_head(token_list, H, W, sd, pre, HEAD_ACTIVATIONS[pre])
        preds[name], preds[name + "_conf"] = p, c
    return preds
