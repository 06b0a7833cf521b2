"""Secondary seams of the hot path: `Block.forward(x, pos)` (src/models/layers/block.py:72-92) and
`Attention.forward(x, pos)` (src/models/layers/attention.py:48-69) over the micro-ops of libwm_b200 — the same kernels
`wm_forward` runs, one block at a time, so that a block of the reference can be swapped and compared on its own.

Same constructor keywords, same `state_dict` keys on the way in (`norm1.*`, `attn.qkv.*`, `attn.q_norm.*`, `attn.k_norm.*`,
`attn.proj.*`, `ls1.gamma`, `norm2.*`, `mlp.fc1.*`, `mlp.fc2.*`, `ls2.gamma`), same shapes in and out.  Built for the
aggregator's geometry: dim 1024, 16 heads of 64, mlp_ratio 4.  The fused QKV kernel derives the 2D-RoPE position of a token
from its index (reference defect D8: no position tensor is read, no host sync per block), so `pos` must be the
aggregator's own layout — `patch_start` tokens at (0, 0), then the row-major (1 + y, 1 + x) grid
(visual_transformer.py:302-306) — which is verified on every call (one host read); anything else raises.
Inference only (eval: dropout and drop-path are identities).  No CPU path.
"""
from __future__ import annotations

import ctypes as C
from typing import Dict, Optional, Tuple

import torch

from . import capi

DIM, HEADS, HEAD_DIM, HIDDEN = 1024, 16, 64, 4096


def _st() -> int:
    return torch.cuda.current_stream().cuda_stream


def _p(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class B200RotaryPositionEmbedding2D:
    """Stands where the reference passes `RotaryPositionEmbedding2D(frequency=100)` (rope.py:80-181).  Inside a block the
    rotation happens in the QKV kernel's epilogue; called on its own (`rope(tokens, positions)`, rope.py:148) it runs the
    stand-alone operator `wm_rope2d`, which takes arbitrary positions.  The tables come from `wm_rope_tables`."""

    def __init__(self, frequency: float = 100.0, scaling_factor: float = 1.0):
        if frequency != 100.0 or scaling_factor != 1.0:
            raise NotImplementedError("the kernel's tables are built for frequency 100, scaling 1 (visual_transformer.py:120)")
        self._tables: Dict[Tuple[int, torch.device], Tuple[torch.Tensor, torch.Tensor]] = {}

    def tables(self, max_pos: int, dev: torch.device) -> Tuple[torch.Tensor, torch.Tensor]:
        key = (max_pos, dev)
        if key not in self._tables:
            cos, sin = (C.c_float * ((max_pos + 1) * 16))(), (C.c_float * ((max_pos + 1) * 16))()
            capi.lib().wm_rope_tables(max_pos, cos, sin)
            self._tables[key] = (torch.tensor(list(cos), device=dev), torch.tensor(list(sin), device=dev))
        return self._tables[key]


    @torch.no_grad()
    def forward(self, tokens: torch.Tensor, positions: torch.Tensor) -> torch.Tensor:
        """tokens [B', heads, L, 64], positions [B', L, 2] (y, x) -> rotated tokens, fp32 (rope.py:148-181; same assertions)."""
        assert tokens.size(-1) % 2 == 0, "Feature dimension must be even"
        assert positions.ndim == 3 and positions.shape[-1] == 2, "Positions must have shape (batch_size, n_tokens, 2)"
        if tokens.ndim != 4 or tokens.size(-1) != HEAD_DIM or not tokens.is_cuda:
            raise ValueError(f"B200RotaryPositionEmbedding2D: tokens must be a CUDA tensor [B, heads, L, {HEAD_DIM}], got "
                             f"{tuple(tokens.shape)} on {tokens.device}")
        Bn, Hh, L, _ = tokens.shape
        if positions.shape[0] != Bn or positions.shape[1] != L:
            raise ValueError(f"positions {tuple(positions.shape)} do not match tokens {tuple(tokens.shape)}")
        t = tokens.to(torch.float32).contiguous()
        pos = positions.to(tokens.device, torch.int64).contiguous()
        max_pos = max(int(pos.max().item()), 0)  # the reference sizes its tables the same way (rope.py:171)
        cs, sn = self.tables(max(512, max_pos), tokens.device)
        out = torch.empty_like(t)
        capi.check(capi.lib().wm_rope2d(_st(), t.data_ptr(), pos.data_ptr(), Bn, Hh, L, max(512, max_pos), cs.data_ptr(), sn.data_ptr(),
                                        out.data_ptr()))
        return out

    __call__ = forward


def grid_of_positions(pos: torch.Tensor) -> Tuple[int, int]:
    """pos [B', L, 2] int64 -> (patch_start, grid_w) of the aggregator's layout; raises if pos is anything else."""
    if pos.ndim != 3 or pos.shape[-1] != 2:
        raise ValueError(f"pos must be [B, L, 2], got {tuple(pos.shape)}")
    L = pos.shape[1]
    gh, gw = int(pos[0, :, 0].max().item()), int(pos[0, :, 1].max().item())
    ps = L - gh * gw
    ok = ps >= 0 and (gh > 0) == (gw > 0)
    if ok:
        i = torch.arange(gh * gw, device=pos.device)
        want = torch.zeros(L, 2, dtype=pos.dtype, device=pos.device)
        if gh * gw:
            want[ps:, 0], want[ps:, 1] = 1 + i // gw, 1 + i % gw
        ok = bool((pos == want[None]).all().item())
    if not ok:
        raise NotImplementedError("B200Block derives RoPE positions from the token index: pos must be `patch_start` tokens at "
                                  "(0, 0) followed by the row-major (1 + y, 1 + x) grid (visual_transformer.py:302-306)")
    return ps, max(gw, 1)


class B200Attention(torch.nn.Module):
    """`Attention` (attention.py:20-69): qkv -> q_norm / k_norm -> rope -> scaled_dot_product_attention -> proj."""

    def __init__(self, dim: int = DIM, num_heads: int = HEADS, qkv_bias: bool = True, proj_bias: bool = True, attn_drop: float = 0.0,
                 proj_drop: float = 0.0, norm_layer=torch.nn.LayerNorm, qk_norm: bool = False, fused_attn: bool = True, rope=None):
        super().__init__()
        if dim != DIM or num_heads != HEADS or getattr(norm_layer, "func", norm_layer) is not torch.nn.LayerNorm:
            raise NotImplementedError("B200Attention is built for dim 1024, 16 heads of 64, LayerNorm QK-norm")
        if rope is not None and not isinstance(rope, B200RotaryPositionEmbedding2D):
            rope = B200RotaryPositionEmbedding2D(getattr(rope, "base_frequency", 100.0), getattr(rope, "scaling_factor", 1.0))
        self.num_heads, self.head_dim, self.scale = num_heads, HEAD_DIM, HEAD_DIM ** -0.5
        self.qk_norm, self.rope = qk_norm, rope
        self.qk_eps = 1e-5
        self.w: Dict[str, torch.Tensor] = {}

    def load_state_dict(self, sd, strict: bool = True, device=None, prefix: str = ""):
        capi.lib()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        f32 = lambda k: sd[prefix + k].detach().to(dev, torch.float32).contiguous()  # noqa: E731
        bf = lambda k: sd[prefix + k].detach().to(dev, torch.float32).to(torch.bfloat16).contiguous()  # noqa: E731
        zeros = lambda n: torch.zeros(n, dtype=torch.float32, device=dev)  # noqa: E731
        w = {"qkv.w": bf("qkv.weight"), "qkv.b": f32("qkv.bias") if prefix + "qkv.bias" in sd else zeros(3 * DIM),
             "proj.w": bf("proj.weight"), "proj.b": f32("proj.bias") if prefix + "proj.bias" in sd else zeros(DIM)}
        if self.qk_norm:
            for n in ("q_norm", "k_norm"):
                w[n + ".w"], w[n + ".b"] = f32(n + ".weight"), f32(n + ".bias")
        assert w["qkv.w"].shape == (3 * DIM, DIM) and w["proj.w"].shape == (DIM, DIM)
        self.w = w
        return torch.nn.modules.module._IncompatibleKeys([], [])

    def _geometry(self, pos, L: int) -> Tuple[int, int]:
        if self.rope is None:
            return 0, 1
        if pos is None:
            raise ValueError("rope is set: pos is required (attention.py:54-56)")
        if pos.shape[1] != L:
            raise ValueError(f"pos has {pos.shape[1]} tokens, x has {L}")
        return grid_of_positions(pos)  # one host read per call: this seam is for unit parity, wm_forward has none

    def attend(self, a_bf16: torch.Tensor, B: int, L: int, pos) -> torch.Tensor:
        """a [B*L, 1024] bf16 (the normed input) -> attention output before `proj`, [B*L, 1024] bf16."""
        if not self.w:
            raise RuntimeError("B200Attention: load_state_dict() has not been called")
        lib, st, dev, w = capi.lib(), _st(), a_bf16.device, self.w
        T = B * L
        ps, gw = self._geometry(pos, L)
        cs = sn = None
        if self.rope is not None:
            cs, sn = self.rope.tables(max(512, (L - ps) // gw, gw), dev)
        q = torch.empty(T, DIM, dtype=torch.bfloat16, device=dev)
        kv = torch.empty(T, 2 * DIM, dtype=torch.bfloat16, device=dev)
        capi.check(lib.wm_gemm_qkv(st, a_bf16.data_ptr(), w["qkv.w"].data_ptr(), w["qkv.b"].data_ptr(), _p(w.get("q_norm.w")),
                                   _p(w.get("q_norm.b")), _p(w.get("k_norm.w")), _p(w.get("k_norm.b")), self.qk_eps, _p(cs), _p(sn), L, ps,
                                   gw, 1.0, T, q.data_ptr(), kv.data_ptr()))
        o = torch.empty(T, DIM, dtype=torch.bfloat16, device=dev)
        capi.check(lib.wm_flash_attn(st, q.data_ptr(), DIM, kv.data_ptr(), kv.data_ptr() + 2 * DIM, 2 * DIM, o.data_ptr(), DIM, B, L, L,
                                     HEADS, self.scale, -1))
        return o

    @torch.no_grad()
    def forward(self, x: torch.Tensor, pos=None) -> torch.Tensor:
        """x [B', L, 1024] fp32 (already normed, as `Block` passes it) -> [B', L, 1024] fp32."""
        B, L, Cd = x.shape
        if Cd != DIM or not x.is_cuda:
            raise ValueError(f"B200Attention: x must be a CUDA tensor [B, L, {DIM}], got {tuple(x.shape)} on {x.device}")
        lib, st, T = capi.lib(), _st(), B * L
        xf = x.to(torch.float32).contiguous()
        a = torch.empty(T, DIM, dtype=torch.bfloat16, device=x.device)
        capi.check(lib.wm_pack_bf16(st, xf.data_ptr(), T, DIM, DIM, a.data_ptr()))
        o = self.attend(a, B, L, pos)
        out = torch.empty(T, DIM, dtype=torch.float32, device=x.device)
        capi.check(lib.wm_gemm_ls_residual(st, o.data_ptr(), DIM, self.w["proj.w"].data_ptr(), DIM, self.w["proj.b"].data_ptr(), None, T,
                                           DIM, DIM, out.data_ptr(), DIM))
        return out.reshape(B, L, DIM)


class B200Block(torch.nn.Module):
    """`Block` (block.py:21-92), eval path: x + ls1(attn(norm1(x), pos)); x + ls2(mlp(norm2(x))).  Seven launches:
    LayerNorm, QKV (+ QK-norm + RoPE), attention, proj (+ LayerScale + residual), LayerNorm, fc1 (+ GELU), fc2 (+ LayerScale +
    residual).  `norm_eps`: 1e-5 for the aggregator's blocks, 1e-6 for DINOv2's (vision_transformer.py:89)."""

    def __init__(self, dim: int = DIM, num_heads: int = HEADS, mlp_ratio: float = 4.0, qkv_bias: bool = True, proj_bias: bool = True,
                 ffn_bias: bool = True, drop: float = 0.0, attn_drop: float = 0.0, init_values=None, drop_path: float = 0.0,
                 act_layer=torch.nn.GELU, norm_layer=torch.nn.LayerNorm, attn_class=None, ffn_layer=None, qk_norm: bool = False,
                 fused_attn: bool = True, rope=None, norm_eps: float = 1e-5):
        super().__init__()
        if dim != DIM or num_heads != HEADS or int(dim * mlp_ratio) != HIDDEN or act_layer is not torch.nn.GELU:
            raise NotImplementedError("B200Block is built for dim 1024, 16 heads, mlp_ratio 4, exact-erf GELU")
        self.attn = B200Attention(dim, num_heads, qkv_bias, proj_bias, qk_norm=qk_norm, rope=rope)
        # the reference passes eps through norm_layer=partial(nn.LayerNorm, eps=1e-6) (vision_transformer.py:89)
        self.init_values, self.norm_eps = init_values, getattr(norm_layer, "keywords", {}).get("eps", norm_eps)
        self.w: Dict[str, torch.Tensor] = {}

    def load_state_dict(self, sd, strict: bool = True, device=None, prefix: str = ""):
        capi.lib()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        self.attn.load_state_dict(sd, device=dev, prefix=prefix + "attn.")
        f32 = lambda k: sd[prefix + k].detach().to(dev, torch.float32).contiguous()  # noqa: E731
        bf = lambda k: sd[prefix + k].detach().to(dev, torch.float32).to(torch.bfloat16).contiguous()  # noqa: E731
        opt = lambda k, n, v: f32(k) if prefix + k in sd else torch.full((n,), v, dtype=torch.float32, device=dev)  # noqa: E731
        self.w = {"n1.w": f32("norm1.weight"), "n1.b": f32("norm1.bias"), "n2.w": f32("norm2.weight"), "n2.b": f32("norm2.bias"),
                  "g1": opt("ls1.gamma", DIM, 1.0), "g2": opt("ls2.gamma", DIM, 1.0),  # nn.Identity when init_values is None
                  "fc1.w": bf("mlp.fc1.weight"), "fc1.b": opt("mlp.fc1.bias", HIDDEN, 0.0),
                  "fc2.w": bf("mlp.fc2.weight"), "fc2.b": opt("mlp.fc2.bias", DIM, 0.0)}
        assert self.w["fc1.w"].shape == (HIDDEN, DIM) and self.w["fc2.w"].shape == (DIM, HIDDEN)
        return torch.nn.modules.module._IncompatibleKeys([], [])

    @torch.no_grad()
    def forward(self, x: torch.Tensor, pos=None) -> torch.Tensor:
        """x [B', L, 1024] fp32, pos [B', L, 2] int64 (or None without rope) -> [B', L, 1024] fp32 (a new tensor)."""
        if not self.w:
            raise RuntimeError("B200Block: load_state_dict() has not been called")
        B, L, Cd = x.shape
        if Cd != DIM or not x.is_cuda:
            raise ValueError(f"B200Block: x must be a CUDA tensor [B, L, {DIM}], got {tuple(x.shape)} on {x.device}")
        lib, st, w, dev, T = capi.lib(), _st(), self.w, x.device, B * L
        xr = x.to(torch.float32).reshape(T, DIM).clone()  # the fp32 residual stream, updated in place by the two epilogues
        a = torch.empty(T, DIM, dtype=torch.bfloat16, device=dev)
        capi.check(lib.wm_layernorm(st, xr.data_ptr(), w["n1.w"].data_ptr(), w["n1.b"].data_ptr(), self.norm_eps, T, a.data_ptr(), 0))
        o = self.attn.attend(a, B, L, pos)
        aw = self.attn.w
        capi.check(lib.wm_gemm_ls_residual(st, o.data_ptr(), DIM, aw["proj.w"].data_ptr(), DIM, aw["proj.b"].data_ptr(), w["g1"].data_ptr(),
                                           T, DIM, DIM, xr.data_ptr(), DIM))
        capi.check(lib.wm_layernorm(st, xr.data_ptr(), w["n2.w"].data_ptr(), w["n2.b"].data_ptr(), self.norm_eps, T, a.data_ptr(), 0))
        h = torch.empty(T, HIDDEN, dtype=torch.bfloat16, device=dev)
        capi.check(lib.wm_gemm_bias_act(st, a.data_ptr(), DIM, w["fc1.w"].data_ptr(), DIM, w["fc1.b"].data_ptr(), T, HIDDEN, DIM,
                                        h.data_ptr(), HIDDEN, 1))
        capi.check(lib.wm_gemm_ls_residual(st, h.data_ptr(), HIDDEN, w["fc2.w"].data_ptr(), HIDDEN, w["fc2.b"].data_ptr(),
                                           w["g2"].data_ptr(), T, DIM, HIDDEN, xr.data_ptr(), DIM))
        return xr.reshape(B, L, DIM)
