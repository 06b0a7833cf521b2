"""CPU oracle for the WorldMirror aggregator hot path.  TEST INFRASTRUCTURE ONLY.

A from-scratch fp32 restatement (plain torch tensor ops on CPU, no nn.Module, no reference
import) of what the reference computes on SURVEY.md §8(a) rows a1-a17.  Only `tests/`,
`__graft_entry__.smoke()` and `bench.py`'s cpu_baseline / `--impl reference` legs may import
this file; the product path (hunyuanworld-mirror_b200/) never does and fails loudly without
its CUDA library.

Parity status: PINNED.  The reference ships no tests or golden vectors for this path
(SURVEY.md §4), so the pin is the reference itself: `tests/golden/make_golden.py` imports
the real `VisualGeometryTransformer` from /root/reference in the build container, loads the
same synthetic state_dict, and stores its outputs under tests/golden/; `tests/test_oracle.py`
checks this file against those fixtures (and, where /root/reference is present, against the
live reference modules).

Every function cites the reference lines it restates (paths relative to /root/reference).
`emulate_bf16=True` rounds at the points where the B200 path holds bf16 (GEMM operands and
outputs, attention operands), which the kernel-level tests use for tight comparisons; the
default is the pure fp32 oracle that the stated tolerances refer to.
"""
from __future__ import annotations

import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

Tensor = torch.Tensor

EMBED = 1024
HEADS = 16
HEAD_DIM = 64
PATCH = 14
NUM_REG = 4
PATCH_START = 7  # cam + 4 reg + pose + ray (visual_transformer.py:99-101)
RESNET_MEAN = (0.485, 0.456, 0.406)  # visual_transformer.py:16-17
RESNET_STD = (0.229, 0.224, 0.225)
ROPE_BASE = 100.0  # visual_transformer.py:62
INTERMEDIATE = (4, 11, 17, 23)  # visual_transformer.py:69


def _r(x: Tensor, emulate_bf16: bool) -> Tensor:
    return x.to(torch.bfloat16).to(torch.float32) if emulate_bf16 else x


# --------------------------------------------------------------------------------------
# Layers
# --------------------------------------------------------------------------------------
def layer_norm(x: Tensor, w: Tensor, b: Tensor, eps: float) -> Tensor:
    """nn.LayerNorm over the last dim (block.py:44,61; attention.py:41-42): biased variance."""
    mu = x.mean(dim=-1, keepdim=True)
    var = ((x - mu) ** 2).mean(dim=-1, keepdim=True)
    return (x - mu) * torch.rsqrt(var + eps) * w + b


def linear(x: Tensor, w: Tensor, b: Optional[Tensor], emulate_bf16: bool = False) -> Tensor:
    """nn.Linear; under emulate_bf16 operands and result are bf16-rounded, accumulation fp32."""
    y = _r(x, emulate_bf16) @ _r(w, emulate_bf16).t()
    if b is not None:
        y = y + b
    return _r(y, emulate_bf16)


def gelu_erf(x: Tensor) -> Tensor:
    """nn.GELU() default = exact erf form (mlp.py:16,25)."""
    return 0.5 * x * (1.0 + torch.erf(x * (1.0 / math.sqrt(2.0))))


def rope_tables(max_pos: int, dim: int = HEAD_DIM // 2, base: float = ROPE_BASE) -> Tuple[Tensor, Tensor]:
    """cos/sin tables [max_pos, dim//2] (rope.py:94-111).  The reference stores cat(ang, ang);
    both halves are equal so we keep one.  inv_freq_i = base^(-2i/dim), angle = pos * inv_freq."""
    expo = torch.arange(0, dim, 2, dtype=torch.float32) / dim
    inv_freq = 1.0 / (base ** expo)
    pos = torch.arange(max_pos, dtype=torch.float32)
    ang = pos[:, None] * inv_freq[None, :]
    return ang.cos(), ang.sin()


def rope_2d(t: Tensor, pos: Tensor) -> Tensor:
    """2D RoPE (rope.py:148-181).  t: [B, H, L, 64]; pos: [B, L, 2] int (y, x).
    Dims [0,32) rotate by y, [32,64) by x; inside each 32-half the pair is (i, i+16)
    (rotate-half, rope.py:123-125): out[i] = t[i]*cos - t[i+16]*sin,
    out[i+16] = t[i+16]*cos + t[i]*sin."""
    half = t.shape[-1] // 2  # 32
    quarter = half // 2  # 16
    cos_t, sin_t = rope_tables(int(pos.max()) + 1, half)
    out = torch.empty_like(t)
    for axis in range(2):
        p = pos[..., axis].long()  # [B, L]
        c = cos_t[p][:, None]  # [B,1,L,16]
        s = sin_t[p][:, None]
        lo = t[..., axis * half: axis * half + quarter]
        hi = t[..., axis * half + quarter: (axis + 1) * half]
        out[..., axis * half: axis * half + quarter] = lo * c - hi * s
        out[..., axis * half + quarter: (axis + 1) * half] = hi * c + lo * s
    return out


def sdpa(q: Tensor, k: Tensor, v: Tensor, chunk: int = 2048) -> Tensor:
    """softmax(q k^T / sqrt(d)) v, no mask (attention.py:59).  [B,H,L,D]; q chunked for memory."""
    scale = q.shape[-1] ** -0.5
    outs = []
    for s in range(0, q.shape[2], chunk):
        a = torch.matmul(q[:, :, s:s + chunk] * scale, k.transpose(-1, -2))
        outs.append(torch.matmul(torch.softmax(a, dim=-1), v))
    return torch.cat(outs, dim=2)


def attention(x: Tensor, pos: Optional[Tensor], sd: Dict[str, Tensor], pre: str,
              qk_norm: bool, emulate_bf16: bool = False) -> Tensor:
    """Attention.forward (attention.py:48-69).  x: [B, L, C]."""
    B, L, C = x.shape
    qkv = linear(x, sd[pre + "qkv.weight"], sd[pre + "qkv.bias"])  # bf16 rounding applied below
    if emulate_bf16:
        # the B200 epilogue keeps the fp32 accumulator through QK-norm and RoPE and rounds once
        qkv = _r(x, True) @ _r(sd[pre + "qkv.weight"], True).t() + sd[pre + "qkv.bias"]
    qkv = qkv.reshape(B, L, 3, HEADS, HEAD_DIM).permute(2, 0, 3, 1, 4)
    q, k, v = qkv[0], qkv[1], qkv[2]
    if qk_norm:
        q = layer_norm(q, sd[pre + "q_norm.weight"], sd[pre + "q_norm.bias"], 1e-5)
        k = layer_norm(k, sd[pre + "k_norm.weight"], sd[pre + "k_norm.bias"], 1e-5)
    if pos is not None:
        q = rope_2d(q, pos)
        k = rope_2d(k, pos)
    q, k, v = _r(q, emulate_bf16), _r(k, emulate_bf16), _r(v, emulate_bf16)
    o = sdpa(q, k, v)
    o = _r(o, emulate_bf16).transpose(1, 2).reshape(B, L, C)
    return linear(o, sd[pre + "proj.weight"], sd[pre + "proj.bias"], emulate_bf16)


def mlp(x: Tensor, sd: Dict[str, Tensor], pre: str, emulate_bf16: bool = False) -> Tensor:
    """Mlp.forward (mlp.py:29-35): fc1 -> erf GELU -> fc2."""
    if emulate_bf16:
        h = _r(x, True) @ _r(sd[pre + "fc1.weight"], True).t() + sd[pre + "fc1.bias"]
        h = _r(gelu_erf(h), True)  # fused bias+GELU epilogue rounds once
    else:
        h = gelu_erf(linear(x, sd[pre + "fc1.weight"], sd[pre + "fc1.bias"]))
    return linear(h, sd[pre + "fc2.weight"], sd[pre + "fc2.bias"], emulate_bf16)


def block(x: Tensor, pos: Optional[Tensor], sd: Dict[str, Tensor], pre: str, eps: float,
          qk_norm: bool, emulate_bf16: bool = False) -> Tensor:
    """Block.forward eval branch (block.py:91-92) with LayerScale (layer_scale.py:17)."""
    n1 = _r(layer_norm(x, sd[pre + "norm1.weight"], sd[pre + "norm1.bias"], eps), emulate_bf16)
    x = x + sd[pre + "ls1.gamma"] * attention(n1, pos, sd, pre + "attn.", qk_norm, emulate_bf16)
    n2 = _r(layer_norm(x, sd[pre + "norm2.weight"], sd[pre + "norm2.bias"], eps), emulate_bf16)
    x = x + sd[pre + "ls2.gamma"] * mlp(n2, sd, pre + "mlp.", emulate_bf16)
    return x


# --------------------------------------------------------------------------------------
# DINOv2 front (vision_transformer.py)
# --------------------------------------------------------------------------------------
def dino_pos_embed(sd: Dict[str, Tensor], gh: int, gw: int) -> Tensor:
    """interpolate_pos_encoding (vision_transformer.py:175-207) with interpolate_offset=0.0,
    antialias=True (visual_transformer.py:117-118).  Returns [1 + gh*gw, C].

    The reference calls it with (w, h) = x.shape[-2:], i.e. its `w` is the image HEIGHT; the
    resize target is therefore (H//14, W//14) = (gh, gw) in (rows, cols) order.  Used verbatim
    when gh*gw == 37*37 and H == W."""
    pe = sd["patch_embed.pos_embed"][0].float()
    n = pe.shape[0] - 1
    m = int(math.isqrt(n))
    if gh * gw == n and gh == gw:
        return pe
    grid = pe[1:].reshape(1, m, m, -1).permute(0, 3, 1, 2)
    grid = F.interpolate(grid, size=(gh, gw), mode="bicubic", antialias=True)
    grid = grid.permute(0, 2, 3, 1).reshape(gh * gw, -1)
    return torch.cat([pe[:1], grid], dim=0)


def patchify(images: Tensor) -> Tensor:
    """[N,3,H,W] -> [N, gh*gw, 588] rows ordered (c, ky, kx) to match conv weight.flatten(1)."""
    N, Cc, H, W = images.shape
    gh, gw = H // PATCH, W // PATCH
    x = images.reshape(N, Cc, gh, PATCH, gw, PATCH).permute(0, 2, 4, 1, 3, 5)
    return x.reshape(N, gh * gw, Cc * PATCH * PATCH)


def dino_forward(images_norm: Tensor, sd: Dict[str, Tensor], dino_depth: int = 24,
                 emulate_bf16: bool = False) -> Tensor:
    """DinoVisionTransformer.forward_features -> x_norm_patchtokens
    (vision_transformer.py:209-221,247-266; patch_embed.py:63-76).  [N,3,H,W] -> [N, g, C]."""
    pre = "patch_embed."
    N, _, H, W = images_norm.shape
    gh, gw = H // PATCH, W // PATCH
    wconv = sd[pre + "patch_embed.proj.weight"].reshape(EMBED, -1)
    tok = linear(patchify(images_norm), wconv, sd[pre + "patch_embed.proj.bias"], emulate_bf16)
    cls = sd[pre + "cls_token"].expand(N, -1, -1)
    x = torch.cat([cls, tok], dim=1) + dino_pos_embed(sd, gh, gw)[None]
    reg = sd[pre + "register_tokens"].expand(N, -1, -1)
    x = torch.cat([x[:, :1], reg, x[:, 1:]], dim=1)  # [N, 1+4+g, C]
    for i in range(dino_depth):
        x = block(x, None, sd, f"{pre}blocks.{i}.", 1e-6, qk_norm=False, emulate_bf16=emulate_bf16)
    x = layer_norm(x, sd[pre + "norm.weight"], sd[pre + "norm.bias"], 1e-6)
    return x[:, 1 + NUM_REG:]


# --------------------------------------------------------------------------------------
# Prior tokens and token assembly (visual_transformer.py:285-306, 343-371, 397-416)
# --------------------------------------------------------------------------------------
def special_tokens(tok: Tensor, n_views: int) -> Tensor:
    """expand_and_flatten_special_tokens (visual_transformer.py:397-416), B=1:
    variant 0 for view 0, variant 1 for every other view.  tok: [1,2,X,C] -> [N,X,C]."""
    return torch.cat([tok[0, 0:1], tok[0, 1:2].expand(n_views - 1, -1, -1)], dim=0)


def token_mlp(v: Tensor, sd: Dict[str, Tensor], pre: str, emulate_bf16: bool = False) -> Tensor:
    """pose_embed / ray_embed: Linear -> SiLU -> Linear (visual_transformer.py:178-204)."""
    h = linear(v, sd[pre + "0.weight"], sd[pre + "0.bias"])
    h = h * torch.sigmoid(h)
    return linear(_r(h, emulate_bf16), sd[pre + "2.weight"], sd[pre + "2.bias"], emulate_bf16)


def depth_tokens(depth: Tensor, sd: Dict[str, Tensor], emulate_bf16: bool = False) -> Tensor:
    """depth_embed = PatchEmbed_Mlp (patch_embed.py:79-93): pixel_unshuffle(14) puts the
    14x14 patch on the channel axis in (ky, kx) order, then Mlp(196 -> 4096 -> 1024).
    depth: [N,H,W] -> [N, g, C]."""
    N, H, W = depth.shape
    gh, gw = H // PATCH, W // PATCH
    x = depth.reshape(N, gh, PATCH, gw, PATCH).permute(0, 1, 3, 2, 4).reshape(N, gh * gw, PATCH * PATCH)
    return mlp(x, sd, "depth_embed.proj.2.", emulate_bf16)


def positions(n_views: int, gh: int, gw: int) -> Tensor:
    """[N, P, 2] int64: (0,0) for the 7 special tokens, (y+1, x+1) row-major for patches
    (rope.py:33-53; visual_transformer.py:299-306)."""
    yy, xx = torch.meshgrid(torch.arange(gh), torch.arange(gw), indexing="ij")
    p = torch.stack([yy.reshape(-1), xx.reshape(-1)], dim=-1) + 1
    p = torch.cat([torch.zeros(PATCH_START, 2, dtype=p.dtype), p], dim=0)
    return p[None].expand(n_views, -1, -1).contiguous()


def assemble_tokens(patch_tokens: Tensor, sd: Dict[str, Tensor], priors, cond_flags: Sequence[int],
                    emulate_bf16: bool = False) -> Tensor:
    """[cam, reg x4, pose, ray, patches(+depth)] (visual_transformer.py:285-295,343-371).
    priors = (depth[1,N,H,W] | None, rays[1,N,4] | None, poses[1,N,7] | None)."""
    N, g, C = patch_tokens.shape
    depth, rays, poses = priors if priors is not None else (None, None, None)
    cam = special_tokens(sd["cam_token"], N)
    reg = special_tokens(sd["reg_token"], N)
    pose_tok = torch.zeros(N, 1, C)
    ray_tok = torch.zeros(N, 1, C)
    if cond_flags[0] == 1 and poses is not None:
        pose_tok = token_mlp(poses.reshape(N, -1).float(), sd, "pose_embed.", emulate_bf16)[:, None]
    if cond_flags[1] == 1 and depth is not None:
        patch_tokens = patch_tokens + depth_tokens(depth.reshape(N, *depth.shape[-2:]).float(), sd, emulate_bf16)
    if cond_flags[2] == 1 and rays is not None:
        ray_tok = token_mlp(rays.reshape(N, -1).float(), sd, "ray_embed.", emulate_bf16)[:, None]
    return torch.cat([cam, reg, pose_tok, ray_tok, patch_tokens], dim=1)


# --------------------------------------------------------------------------------------
# The aggregator (visual_transformer.py:250-341)
# --------------------------------------------------------------------------------------
def aggregator_stack(x: Tensor, pos: Tensor, sd: Dict[str, Tensor], depth: int = 24,
                     intermediate: Sequence[int] = INTERMEDIATE, emulate_bf16: bool = False) -> List[Tensor]:
    """24 x (frame Block on [N,P,C], global Block on [1,N*P,C]) (visual_transformer.py:309-339).
    x: [N,P,C], pos: [N,P,2].  Returns the [1,N,P,2C] tensors at `intermediate` layers."""
    N, P, C = x.shape
    outs = []
    for i in range(depth):
        x = block(x, pos, sd, f"frame_blocks.{i}.", 1e-5, True, emulate_bf16)
        local = x
        x = block(x.reshape(1, N * P, C), pos.reshape(1, N * P, 2), sd, f"global_blocks.{i}.", 1e-5, True,
                  emulate_bf16).reshape(N, P, C)
        if i in intermediate:
            outs.append(torch.cat([local, x], dim=-1)[None])
    return outs


def aggregator_forward(images: Tensor, sd: Dict[str, Tensor], priors=None, cond_flags: Sequence[int] = (0, 0, 0),
                       depth: int = 24, dino_depth: int = 24, intermediate: Sequence[int] = INTERMEDIATE,
                       emulate_bf16: bool = False, return_stages: bool = False):
    """VisualGeometryTransformer.forward (visual_transformer.py:250-341), B=1, enable_cond=True.
    images: [1,N,3,H,W] in [0,1].  Returns (list of [1,N,P,2048] fp32, patch_start_idx=7)."""
    b, N, ch, H, W = images.shape
    if ch != 3:
        raise ValueError(f"Expected 3 input channels, got {ch}")  # visual_transformer.py:272-273
    assert b == 1
    assert H % PATCH == 0 and W % PATCH == 0  # patch_embed.py:67-68
    mean = torch.tensor(RESNET_MEAN).reshape(1, 3, 1, 1)
    std = torch.tensor(RESNET_STD).reshape(1, 3, 1, 1)
    imgs = (images[0].float() - mean) / std
    patch_tokens = dino_forward(imgs, sd, dino_depth, emulate_bf16)
    x = assemble_tokens(patch_tokens, sd, priors, cond_flags, emulate_bf16)
    pos = positions(N, H // PATCH, W // PATCH)
    outs = aggregator_stack(x, pos, sd, depth, intermediate, emulate_bf16)
    if return_stages:
        return outs, PATCH_START, {"dino_patch_tokens": patch_tokens, "tokens0": x}
    return outs, PATCH_START
