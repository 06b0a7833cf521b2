"""B200 prediction heads behind the reference's module interface.

`B200DPTHead` mirrors `DPTHead` (src/models/heads/dense_head.py:11-260, is_gsdpt=False): same constructor keywords,
same `state_dict` keys on the way in, same `forward(token_list, images, patch_start_idx, frames_chunk_size)` and the
same return value `(pred [B,S,H,W,C-1], conf [B,S,H,W])` in fp32.  Every convolution runs as an implicit GEMM on
tcgen05 (csrc/conv.cu) over NHWC fp16 activations (10-bit mantissa, as TF32) with fp32 accumulation and fp32 epilogue
arithmetic (the reference runs its heads in fp32, worldmirror.py:146; bf16 activations measured 1.5e-2 on pts3d, a
10-bit mantissa 1.2e-3); bias, ReLU, the residual adds of
ResidualConvUnit / FeatureFusionBlock, the positional embedding and the output activations are fused into the
epilogues; LayerNorm, bilinear resampling and the ConvTranspose scatter are bandwidth kernels (csrc/headops.cu).
Two algebraic rewrites, both exact in real arithmetic:
  * FeatureFusionBlock runs `out_conv` (1x1) BEFORE the bilinear upsample instead of after (dense_head.py:529-530):
    both are linear and the interpolation weights sum to one, so they commute; 4x fewer FLOPs.
  * nn.ReLU(inplace=True) inside ResidualConvUnit overwrites its input (dense_head.py:448), so the skip connection
    adds relu(x); producers therefore store relu(x) only.
PyTorch is used for device memory and for packing weights at load time.  There is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import math
import os
from typing import Dict, List, Sequence, Tuple

import torch

from . import capi

PATCH = 14
ACT = torch.float16  # storage type of every activation map and convolution weight of the DPT heads (csrc/conv.cu)
_ACT = {"exp": 0, "inv_log": 1, "norm": 2}
LAUNCHES = [0]  # kernels launched by the heads since the caller last reset it (bench.py: gpu_launches)


def _k(status: int) -> None:
    """Check the status of one kernel-launching library call and count it."""
    capi.check(status)
    LAUNCHES[0] += 1


def _st():
    return torch.cuda.current_stream().cuda_stream


def _p(t):
    return None if t is None else t.data_ptr()


def round_tf32(t: torch.Tensor) -> torch.Tensor:
    """Round fp32 to TF32 (10-bit mantissa), nearest with ties away from zero = PTX cvt.rna.tf32.f32.  Kept for tests
    that want operands exactly representable in the heads' fp16 storage (same mantissa width)."""
    i = t.contiguous().view(torch.int32)
    return ((i + 0x1000) & ~0x1FFF).view(torch.float32)


def uv_tables(w: int, h: int, channels: int, aspect: float, ratio: float = 0.1) -> Tuple[torch.Tensor, torch.Tensor]:
    """Separable form of `_apply_pos_embed` (dense_head.py:262-271; utils/grid.py:4-89): the embedding of pixel (x, y)
    is [sin(u w_k), cos(u w_k), sin(v w_k), cos(v w_k)] * ratio with u = u(x), v = v(y), w_k = 100^(-k/(C/4)) in float64.
    Returns (pos_u [w, C/2], pos_v [h, C/2]) fp32 on the CPU."""
    diag = math.sqrt(aspect * aspect + 1.0)
    sx, sy = aspect / diag, 1.0 / diag
    u = torch.linspace(-sx * (w - 1) / w, sx * (w - 1) / w, w, dtype=torch.float32).double()
    v = torch.linspace(-sy * (h - 1) / h, sy * (h - 1) / h, h, dtype=torch.float32).double()
    omega = 1.0 / 100.0 ** (torch.arange(channels // 4, dtype=torch.float64) / (channels / 4.0))
    au, av = u[:, None] * omega[None], v[:, None] * omega[None]
    pu = torch.cat([au.sin(), au.cos()], dim=1).float() * ratio
    pv = torch.cat([av.sin(), av.cos()], dim=1).float() * ratio
    return pu.contiguous(), pv.contiguous()


_SAT_HOST: Dict[int, torch.Tensor] = {}


def fp16_saturation(device=None, reset: bool = True, sync: bool = True) -> int:
    """Entries of the DPT heads' fp16 maps that left the fp16 range (and were clamped) since the last reset
    (wm_fp16_saturation, csrc/conv.cu).  sync=False only enqueues the copy on the current stream and returns the value
    of the PREVIOUS read (valid once the stream has been synchronised)."""
    dev = torch.device("cuda", torch.cuda.current_device()) if device is None else torch.device(device)
    idx = dev.index if dev.index is not None else torch.cuda.current_device()
    if idx not in _SAT_HOST:
        _SAT_HOST[idx] = torch.zeros(1, dtype=torch.int32).pin_memory()
    host = _SAT_HOST[idx]
    prev = int(host.item())
    with torch.cuda.device(idx):
        st = torch.cuda.current_stream()
        _k(capi.lib().wm_fp16_saturation(st.cuda_stream, host.data_ptr(), int(reset)))
        if sync:
            st.synchronize()
            return int(host.item())
    return prev


class Fp16RangeError(FloatingPointError):
    """Activations of a prediction head exceeded the fp16 range of its maps (they were clamped to +-65504)."""


class B200DPTHead(torch.nn.Module):
    def __init__(self, dim_in: int = 2048, patch_size: int = 14, output_dim: int = 4, activation: str = "inv_log+expp1",
                 features: int = 256, out_channels: Sequence[int] = (256, 512, 1024, 1024), pos_embed: bool = True,
                 down_ratio: int = 1, is_gsdpt: bool = False):
        super().__init__()
        if down_ratio != 1 or not pos_embed or patch_size != PATCH:
            raise NotImplementedError("B200DPTHead covers the configurations WorldMirror constructs (down_ratio 1, pos_embed)")
        if dim_in not in (1024, 2048) or features != 256 or tuple(out_channels) != (256, 512, 1024, 1024):
            raise NotImplementedError("B200DPTHead is built for dim_in 2048, features 256, out_channels (256,512,1024,1024)")
        act, conf = activation.split("+") if "+" in activation else (activation, "expp1")
        if act not in _ACT or conf != "expp1" or output_dim not in (2, 4):
            raise ValueError(f"Unknown attribute activation: {activation}")
        self.dim_in, self.output_dim, self.activation, self.act = dim_in, output_dim, activation, _ACT[act]
        self.is_gsdpt = bool(is_gsdpt)
        self.ln_cache = None  # dict shared by the heads of one forward (B200WorldMirror sets and clears it): normalised token rows
        self.tail_views = int(os.environ.get("WM_DPT_TAIL_VIEWS", "0")) or 1 << 30  # views per pass of the high-resolution tail (0: all)
        self.oc = tuple(out_channels)
        self.w: Dict[str, torch.Tensor] = {}
        self._pos: Dict[tuple, tuple] = {}
        self._loaded = False

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, sd, strict: bool = True, device=None):
        """Reference keys and shapes (DPTHead.state_dict()).  Convolution weights are repacked to fp16
        [Cout, (ky, kx, ci)]; ConvTranspose2d weights [ci, co, ky, kx] to [(ky, kx, co), ci] with the bias tiled."""
        capi.lib()  # raises if the CUDA library is missing: no fallback
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        w = {}

        def conv(key):
            t = sd[key].detach().to(dev, torch.float32)
            co, ci, kh, kw = t.shape
            if float(t.abs().max()) > 65504.0:
                raise Fp16RangeError(f"{key}: |weight| exceeds the fp16 range of the convolution operands")
            return t.permute(0, 2, 3, 1).reshape(co, kh * kw * ci).to(ACT).contiguous()

        def vec(key):
            return sd[key].detach().to(dev, torch.float32).contiguous()

        # LayerNorm's affine part is folded into the 1x1 projections that consume it,
        #   W (x_hat * g + b) + c = (W diag(g)) x_hat + (W b + c),
        # so that the normalised rows x_hat do not depend on the head: the four heads of a forward share them
        # (`ln_cache`, 4 instead of 16 LayerNorm(2048) launches over the same token rows)
        gamma, beta = vec("norm.weight"), vec("norm.bias")
        w["norm.w"], w["norm.b"] = torch.ones_like(gamma), torch.zeros_like(beta)
        for i in range(4):
            Wp = sd[f"projects.{i}.weight"].detach().to(dev, torch.float32)  # [co, 2048, 1, 1]
            folded = Wp * gamma[None, :, None, None]
            if float(folded.abs().max()) > 65504.0:
                raise Fp16RangeError(f"projects.{i}.weight * norm.weight exceeds the fp16 range of the convolution operands")
            w[f"proj{i}.w"] = folded.permute(0, 2, 3, 1).reshape(Wp.shape[0], -1).to(ACT).contiguous()
            w[f"proj{i}.b"] = (vec(f"projects.{i}.bias") + Wp.reshape(Wp.shape[0], -1) @ beta).contiguous()
            w[f"rn{i}.w"] = conv(f"scratch.layer{i + 1}_rn.weight")
        for i, k in ((0, 4), (1, 2)):
            t = sd[f"resize_layers.{i}.weight"].detach().to(dev, torch.float32)  # [ci, co, k, k]
            ci, co = t.shape[0], t.shape[1]
            w[f"up{i}.w"] = t.permute(2, 3, 1, 0).reshape(k * k * co, ci).to(ACT).contiguous()
            w[f"up{i}.b"] = vec(f"resize_layers.{i}.bias").repeat(k * k).contiguous()
        w["down3.w"], w["down3.b"] = conv("resize_layers.3.weight"), vec("resize_layers.3.bias")
        for r in (1, 2, 3, 4):
            p = f"scratch.refinenet{r}."
            w[f"rf{r}.out.w"], w[f"rf{r}.out.b"] = conv(p + "out_conv.weight"), vec(p + "out_conv.bias")
            for u in ((1, 2) if r != 4 else (2,)):
                for c in (1, 2):
                    w[f"rf{r}.u{u}.c{c}.w"] = conv(p + f"resConfUnit{u}.conv{c}.weight")
                    w[f"rf{r}.u{u}.c{c}.b"] = vec(p + f"resConfUnit{u}.conv{c}.bias")
        w["oc1.w"], w["oc1.b"] = conv("scratch.output_conv1.weight"), vec("scratch.output_conv1.bias")
        w["oc2.w"], w["oc2.b"] = conv("scratch.output_conv2.0.weight"), vec("scratch.output_conv2.0.bias")
        w["oc3.w"] = sd["scratch.output_conv2.2.weight"].detach().to(dev, torch.float32).reshape(self.output_dim, 32).contiguous()
        w["oc3.b"] = vec("scratch.output_conv2.2.bias")
        if self.is_gsdpt:  # input_merger = Conv2d(3, 128, 7, 1, 3): [co, ci, ky, kx] -> [co, ky, (kx, ci) padded 21 -> 64]
            t = sd["input_merger.0.weight"].detach().to(dev, torch.float32)
            co = t.shape[0]
            rows = torch.zeros(co, 7, 64, dtype=torch.float32, device=dev)
            rows[:, :, :21] = t.permute(0, 2, 3, 1).reshape(co, 7, 21)
            w["im.w"], w["im.b"] = rows.reshape(co, 7 * 64).to(ACT).contiguous(), vec("input_merger.0.bias")
        assert tuple(w["proj0.w"].shape) == (self.oc[0], self.dim_in)
        self.w, self._loaded, self._dev = w, True, dev
        return torch.nn.modules.module._IncompatibleKeys([], [])

    def _pos_tables(self, gh, gw, H, W):
        key = (gh, gw, H, W)
        if key not in self._pos:
            tabs = []
            for c in self.oc:  # positional table of the projected features, fp32 [gh, gw, c]
                pu, pv = uv_tables(gw, gh, c, W / H)
                full = torch.cat([pu[None].expand(gh, -1, -1), pv[:, None].expand(-1, gw, -1)], dim=-1)
                tabs.append(full.to(self._dev).contiguous())
            pu, pv = uv_tables(W, H, 128, W / H)
            if len(self._pos) >= 4:  # a few live geometries (tables of one geometry: ~14 MB at 518 x 518): callers that alternate
                self._pos.pop(next(iter(self._pos)))  # between shapes do not rebuild them every call; the oldest one goes
            self._pos[key] = (tabs, pu.to(self._dev), pv.to(self._dev))
        return self._pos[key]

    # ------------------------------------------------------------------ ops
    def _conv(self, x, n, H, W, cin, wkey, taps, cout, bias=None, res1=None, bcast=0, res2=None, relu=0):
        out = torch.empty(n, H, W, cout, dtype=ACT, device=x.device)
        _k(capi.lib().wm_conv2d_nhwc(_st(), x.data_ptr(), n, H, W, cin, self.w[wkey].data_ptr(), taps, cout, _p(bias),
                                             _p(res1), bcast, _p(res2), relu, out.data_ptr()))
        return out

    def _resize(self, x, n, Hi, Wi, c, Ho, Wo, pu=None, pv=None):
        out = torch.empty(n, Ho, Wo, c, dtype=ACT, device=x.device)
        _k(capi.lib().wm_bilinear_nhwc(_st(), x.data_ptr(), n, Hi, Wi, c, Ho, Wo, _p(pu), _p(pv), out.data_ptr()))
        return out

    def _rcu(self, r, unit, x, n, H, W, extra=None, relu_out=0):
        """ResidualConvUnit on x = relu(input): conv2(relu(conv1(x))) + x [+ extra], optional ReLU on the sum."""
        w = self.w
        t = self._conv(x, n, H, W, 256, f"rf{r}.u{unit}.c1.w", 9, 256, w[f"rf{r}.u{unit}.c1.b"], relu=1)
        return self._conv(t, n, H, W, 256, f"rf{r}.u{unit}.c2.w", 9, 256, w[f"rf{r}.u{unit}.c2.b"], res1=x, res2=extra, relu=relu_out)

    # ------------------------------------------------------------------ forward
    @torch.no_grad()
    def forward(self, token_list: List[torch.Tensor], images: torch.Tensor, patch_start_idx: int, frames_chunk_size: int = 8):
        if not self._loaded:
            raise RuntimeError("B200DPTHead: load_state_dict() has not been called")
        B, S, _, H, W = images.shape
        if not token_list[0].is_cuda:
            raise RuntimeError("B200DPTHead needs CUDA tensors (there is no CPU path)")
        chunk = S if (frames_chunk_size is None or frames_chunk_size >= S) else frames_chunk_size
        assert chunk > 0
        dev = token_list[0].device
        na = self.output_dim - 1
        pred = torch.empty(B, S, H, W, na, dtype=torch.float32, device=dev)
        conf = torch.empty(B, S, H, W, dtype=torch.float32, device=dev)
        feat = torch.empty(B, S, H, W, 128, dtype=ACT, device=dev) if self.is_gsdpt else None
        if self.is_gsdpt:
            images = images.to(dev, torch.float32).contiguous()
        with torch.cuda.device(dev):
            for b in range(B):
                for f0 in range(0, S, chunk):
                    f1 = min(f0 + chunk, S)
                    self._forward_chunk([t[b, f0:f1] for t in token_list], H, W, patch_start_idx, pred[b, f0:f1], conf[b, f0:f1],
                                        images[b, f0:f1] if self.is_gsdpt else None, feat[b, f0:f1] if self.is_gsdpt else None)
        if self.is_gsdpt:  # (features [B, S, 128, H, W], predictions, confidence) as dense_head.py:232-245; channels-last memory
            return feat.permute(0, 1, 4, 2, 3), pred, conf
        return pred, conf

    def _forward_chunk(self, toks, H, W, psi, pred, conf, images=None, feat=None):
        lib, w = capi.lib(), self.w
        n, P, Cd = toks[0].shape
        gh, gw = H // PATCH, W // PATCH
        g = gh * gw
        assert P == psi + g and Cd == self.dim_in
        tabs, pu, pv = self._pos_tables(gh, gw, H, W)
        dev = toks[0].device
        rn, sizes = [], []
        cache = self.ln_cache
        for i, t in enumerate(toks):
            t = t.to(torch.float32).contiguous()
            key = (i, t.data_ptr(), n, P, int(_st()))
            ln = cache.get(key) if cache is not None else None
            if ln is None:
                ln = torch.empty(n, gh, gw, Cd, dtype=ACT, device=dev)
                _k(lib.wm_layernorm_rows(_st(), t.data_ptr(), w["norm.w"].data_ptr(), w["norm.b"].data_ptr(), 1e-5, n * g, g, P,
                                                 psi, Cd, ln.data_ptr()))
                if cache is not None:
                    cache[key] = ln
            c = self.oc[i]
            f = self._conv(ln, n, gh, gw, Cd, f"proj{i}.w", 1, c, w[f"proj{i}.b"], res1=tabs[i], bcast=1)
            h_, w_ = gh, gw
            if i in (0, 1):  # ConvTranspose2d(c, c, k, stride k) = 1x1 convolution to k*k*c channels + scatter
                k = 4 if i == 0 else 2
                tmp = self._conv(f, n, gh, gw, c, f"up{i}.w", 1, k * k * c, w[f"up{i}.b"])
                f = torch.empty(n, gh * k, gw * k, c, dtype=ACT, device=dev)
                _k(lib.wm_pixel_shuffle(_st(), tmp.data_ptr(), n, gh, gw, c, k, f.data_ptr()))
                h_, w_ = gh * k, gw * k
            elif i == 3:  # Conv2d(c, c, 3, stride 2, padding 1) = im2col + 1x1 convolution over 9c channels
                h_, w_ = (gh - 1) // 2 + 1, (gw - 1) // 2 + 1
                col = torch.empty(n, h_, w_, 9 * c, dtype=ACT, device=dev)
                _k(lib.wm_im2col_s2(_st(), f.data_ptr(), n, gh, gw, c, col.data_ptr()))
                f = self._conv(col, n, h_, w_, 9 * c, "down3.w", 1, c, w["down3.b"])
            # scratch.layer{i}_rn (no bias); only relu(rn) is ever read (in-place ReLU of the consuming RCU)
            rn.append(self._conv(f, n, h_, w_, c, f"rn{i}.w", 9, 256, relu=1))
            sizes.append((h_, w_))
        # refinenet4 .. refinenet2 (dense_head.py:283-291) on all views of the chunk: small maps
        path = None
        for r in (4, 3, 2):
            h_, w_ = sizes[r - 1]
            if r == 4:
                x = rn[3]
            else:  # x = relu(path + RCU1(skip))
                x = self._rcu(r, 1, rn[r - 1], n, h_, w_, extra=path, relu_out=1)
            u = self._rcu(r, 2, x, n, h_, w_)
            v = self._conv(u, n, h_, w_, 256, f"rf{r}.out.w", 1, 256, w[f"rf{r}.out.b"])
            ho, wo = sizes[r - 2]
            path = self._resize(v, n, h_, w_, 256, ho, wo)
        # refinenet1, output_conv1, the resize to the image size and the final convolutions work on maps of 148^2 .. 518^2
        # pixels: 90 .. 550 MB for 8 views, far beyond the 126 MB L2.  They run `tail_views` views at a time so that what
        # one kernel writes is still in L2 when the next one reads it.
        h_, w_ = sizes[0]
        tv = max(1, min(n, int(self.tail_views)))
        for v0 in range(0, n, tv):
            m = min(tv, n - v0)
            x = self._rcu(1, 1, rn[0][v0:v0 + m], m, h_, w_, extra=path[v0:v0 + m], relu_out=1)
            u = self._rcu(1, 2, x, m, h_, w_)
            v = self._conv(u, m, h_, w_, 256, "rf1.out.w", 1, 256, w["rf1.out.b"])
            hp, wp = 2 * h_, 2 * w_
            p1 = self._resize(v, m, h_, w_, 256, hp, wp)
            o1 = self._conv(p1, m, hp, wp, 256, "oc1.w", 9, 128, w["oc1.b"])
            fused = self._resize(o1, m, hp, wp, 128, gh * PATCH, gw * PATCH, pu, pv)
            _k(lib.wm_conv2d_final(_st(), fused.data_ptr(), m, H, W, 128, w["oc2.w"].data_ptr(), w["oc2.b"].data_ptr(),
                                   w["oc3.w"].data_ptr(), w["oc3.b"].data_ptr(), self.output_dim, self.act,
                                   pred[v0:v0 + m].data_ptr(), conf[v0:v0 + m].data_ptr()))
            if self.is_gsdpt:
                # fused + relu(input_merger(images)) (dense_head.py:239-242): the 7 horizontal taps of the 3-channel image
                # become 21 (of 64) channels of an NHWC map, the 7 vertical taps run as an implicit GEMM with ReLU before
                # the residual
                assert H == gh * PATCH and W == gw * PATCH
                rows = torch.empty(m, H, W, 64, dtype=ACT, device=dev)
                _k(lib.wm_im2row7(_st(), images[v0:v0 + m].data_ptr(), m, H, W, rows.data_ptr()))
                _k(lib.wm_conv2d_nhwc(_st(), rows.data_ptr(), m, H, W, 64, w["im.w"].data_ptr(), 7, 128, w["im.b"].data_ptr(),
                                      fused.data_ptr(), 0, None, 2, feat[v0:v0 + m].data_ptr()))


class B200CameraHead(torch.nn.Module):
    """Mirror of `CameraHead` (src/models/heads/camera_head.py:10-104): same constructor keywords, `state_dict` keys and
    `forward(feat_seq, steps=4) -> list of [B, S, 9]` (activated [t(3), quat xyzw(4), fov(2)] of every refinement step).
    fp32 throughout, like the reference (worldmirror.py:146); every linear layer is a weight-streaming skinny GEMM
    (csrc/camhead.cu).  Under sequence parallelism pass the camera tokens of ALL views (the trunk attends across views,
    camera_head.py:73,92): `hunyuanworld-mirror_b200.parallel.gather_outputs`."""

    def __init__(self, dim_in: int = 2048, trunk_depth: int = 4, num_heads: int = 16, mlp_ratio: int = 4, init_values: float = 0.01,
                 trans_act: str = "linear", quat_act: str = "linear", fl_act: str = "relu"):
        super().__init__()
        if (dim_in, num_heads, mlp_ratio) != (2048, 16, 4) or (trans_act, quat_act, fl_act) != ("linear", "linear", "relu"):
            raise NotImplementedError("B200CameraHead is built for WorldMirror's configuration (dim 2048, 16 heads, linear/linear/relu)")
        self.dim, self.depth = dim_in, trunk_depth
        self.w: Dict[str, torch.Tensor] = {}
        self._graphs: Dict[tuple, object] = {}
        self._loaded = False

    def load_state_dict(self, sd, strict: bool = True, device=None):
        capi.lib()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        need = ["init_token", "token_norm.weight", "token_norm.bias", "out_norm.weight", "out_norm.bias", "param_embed.weight",
                "param_embed.bias", "adapt_norm_gen.1.weight", "adapt_norm_gen.1.bias", "param_predictor.fc1.weight",
                "param_predictor.fc1.bias", "param_predictor.fc2.weight", "param_predictor.fc2.bias"]
        for i in range(self.depth):
            need += [f"refine_net.{i}.{k}" for k in ("norm1.weight", "norm1.bias", "attn.qkv.weight", "attn.qkv.bias", "attn.proj.weight",
                                                     "attn.proj.bias", "ls1.gamma", "norm2.weight", "norm2.bias", "mlp.fc1.weight",
                                                     "mlp.fc1.bias", "mlp.fc2.weight", "mlp.fc2.bias", "ls2.gamma")]
        missing = [k for k in need if k not in sd]
        if missing:
            raise RuntimeError(f"Missing key(s) in state_dict: {missing[:4]}...")
        self.w = {k: sd[k].detach().to(dev, torch.float32).contiguous() for k in need}
        self._graphs = {}
        self._loaded = True
        return torch.nn.modules.module._IncompatibleKeys([], [k for k in sd if k not in need])

    def _linear(self, x, key, M, silu_in=0, act=0, resid=None, gamma=None, out=None):
        W, b = self.w[key + ".weight"], self.w[key + ".bias"]
        N, K = W.shape
        if out is None:
            out = torch.empty(M, N, dtype=torch.float32, device=x.device)
        _k(capi.lib().wm_skinny_linear(_st(), x.data_ptr(), K, W.data_ptr(), b.data_ptr(), M, N, K, silu_in, act, _p(resid),
                                               _p(gamma), out.data_ptr(), N))
        return out

    def _ln(self, x, stride, key, eps, M, mod=None):
        y = torch.empty(M, self.dim, dtype=torch.float32, device=x.device)
        w = self.w[key + ".weight"] if key else None
        b = self.w[key + ".bias"] if key else None
        _k(capi.lib().wm_cam_layernorm(_st(), x.data_ptr(), stride, _p(w), _p(b), eps, M, _p(mod), y.data_ptr()))
        return y

    def _run(self, feat, stride, S, steps, outs, shard=None):
        """One sample: `feat` holds the camera tokens, rows `stride` floats apart; outs[step] is a [S, 9] tensor.
        shard = (group, world, rank): the S tokens are THIS RANK'S views of a scene of S * world views.  Every layer of
        the head is per token except the trunk attention (camera_head.py:73,92 attends across the views), so only the
        [S, 6144] q|k|v rows of each block are all-gathered; attention runs over all S * world tokens (a few
        microseconds) and the rank keeps its own rows.  The weights (0.8 GB per refinement step) are streamed once per
        rank for S rows instead of S * world."""
        lib, w, dev, Cd = capi.lib(), self.w, feat.device, self.dim
        if shard is not None:
            import torch.distributed as dist

            group, world, rank = shard
            qkv_all = torch.empty(S * world, 3 * Cd, dtype=torch.float32, device=dev)
            att_all = torch.empty(S * world, Cd, dtype=torch.float32, device=dev)
        tok = self._ln(feat, stride, "token_norm", 1e-5, S)
        pred = torch.empty(S, 9, dtype=torch.float32, device=dev)
        emb = torch.empty(S, Cd, dtype=torch.float32, device=dev)
        for step in range(steps):
            src = w["init_token"] if step == 0 else pred
            _k(lib.wm_cam_embed(_st(), src.data_ptr(), int(step == 0), 9, w["param_embed.weight"].data_ptr(),
                                w["param_embed.bias"].data_ptr(), S, Cd, emb.data_ptr()))
            mod = self._linear(emb, "adapt_norm_gen.1", S, silu_in=1)
            x = self._ln(tok, Cd, None, 1e-6, S, mod=mod)
            for i in range(self.depth):
                p = f"refine_net.{i}."
                n1 = self._ln(x, Cd, p + "norm1", 1e-5, S)
                qkv = self._linear(n1, p + "attn.qkv", S)
                if shard is None:
                    att = torch.empty(S, Cd, dtype=torch.float32, device=dev)
                    _k(lib.wm_cam_attention(_st(), qkv.data_ptr(), S, att.data_ptr()))
                else:
                    dist.all_gather_into_tensor(qkv_all, qkv, group=group)
                    _k(lib.wm_cam_attention(_st(), qkv_all.data_ptr(), S * world, att_all.data_ptr()))
                    att = att_all[rank * S:(rank + 1) * S]
                self._linear(att, p + "attn.proj", S, resid=x, gamma=w[p + "ls1.gamma"], out=x)
                n2 = self._ln(x, Cd, p + "norm2", 1e-5, S)
                hid = self._linear(n2, p + "mlp.fc1", S, act=1)
                self._linear(hid, p + "mlp.fc2", S, resid=x, gamma=w[p + "ls2.gamma"], out=x)
            xo = self._ln(x, Cd, "out_norm", 1e-5, S)
            hid = self._linear(xo, "param_predictor.fc1", S, act=1)
            _k(lib.wm_cam_predict(_st(), hid.data_ptr(), hid.shape[1], w["param_predictor.fc2.weight"].data_ptr(),
                                  w["param_predictor.fc2.bias"].data_ptr(), S, int(step == 0), pred.data_ptr(), outs[step].data_ptr()))

    def _graph_for(self, S, steps, dev, shard=None):
        """The head is ~137 launches of a few microseconds each: captured once per (views, steps) into a CUDA graph so
        that the launch latency of issuing them from Python does not bound it.  WM_CAM_GRAPH=0 keeps eager launches."""
        import os

        key = (S, steps, dev.index, None if shard is None else (shard[1], shard[2]))
        if key in self._graphs:
            return self._graphs[key]
        entry = None
        # sharded: eager (16 small collectives issued from Python, same stream order on all ranks).  Capturing the
        # all-gathers into the graph works and gives the same numbers, but processes that hold a captured NCCL collective
        # did not exit cleanly on the 2-GPU box (torchrun had to kill them), so it is not offered.
        mode = os.environ.get("WM_CAM_GRAPH", "1")
        if mode != "0" and shard is None:
            tok_in = torch.zeros(S, self.dim, dtype=torch.float32, device=dev)
            outs = [torch.zeros(S, 9, dtype=torch.float32, device=dev) for _ in range(steps)]
            self._run(tok_in, self.dim, S, steps, outs, shard)  # warm-up outside capture
            torch.cuda.synchronize(dev)
            before = LAUNCHES[0]
            graph = torch.cuda.CUDAGraph()
            with torch.cuda.graph(graph):
                self._run(tok_in, self.dim, S, steps, outs, shard)
            entry = (graph, tok_in, outs, LAUNCHES[0] - before)
        self._graphs[key] = entry
        return entry

    @torch.no_grad()
    def forward(self, feat_seq: list, steps: int = 4, shard=None) -> list:
        """shard = (group, world, rank): `feat_seq[-1]` holds this rank's views only and the result is this rank's rows
        (see `_run`); every rank must call with the same number of views."""
        if not self._loaded:
            raise RuntimeError("B200CameraHead: load_state_dict() has not been called")
        latest = feat_seq[-1]
        if not latest.is_cuda:
            raise RuntimeError("B200CameraHead needs CUDA tensors (there is no CPU path)")
        B, S, P, Cd = latest.shape
        assert Cd == self.dim
        dev = latest.device
        outs = [torch.empty(B, S, 9, dtype=torch.float32, device=dev) for _ in range(steps)]
        with torch.cuda.device(dev):
            g = self._graph_for(S, steps, dev, shard)
            for bi in range(B):
                feat = latest[bi].to(torch.float32)
                if g is not None:
                    graph, tok_in, gouts, n_launch = g
                    tok_in.copy_(feat[:, 0])  # camera token = row 0 of every view
                    graph.replay()
                    LAUNCHES[0] += n_launch
                    for step in range(steps):
                        outs[step][bi].copy_(gouts[step])
                else:
                    if not feat.is_contiguous():
                        feat = feat.contiguous()
                    self._run(feat, P * Cd, S, steps, [o[bi] for o in outs], shard)
        return outs


def camera_matrices(camera_params: torch.Tensor, H: int, W: int):
    """camera_params [S, 9] -> (camera-to-world [S, 4, 4], K [S, 3, 3]): `vector_to_camera_matrices` followed by the
    inverse of the homogeneous world-to-camera matrix, as `WorldMirror._gen_all_preds` (worldmirror.py:161-175) and
    `prepare_prediction_cameras` (rasterization.py:506-530) do.  One kernel (csrc/splat.cu: splat_camera_kernel); the
    inverse is the closed form [R^T | -R^T t] instead of the reference's LU-based torch.linalg.inv."""
    S, dev = camera_params.shape[0], camera_params.device
    cam = torch.empty(S, 16, dtype=torch.float32, device=dev)
    cp = camera_params.to(torch.float32).contiguous()
    _k(capi.lib().wm_splat_cameras(_st(), cp.data_ptr(), S, H, W, cam.data_ptr()))
    c2w = torch.zeros(S, 4, 4, dtype=torch.float32, device=dev)
    c2w[:, :3, :3], c2w[:, :3, 3], c2w[:, 3, 3] = cam[:, :9].reshape(S, 3, 3), cam[:, 9:12], 1.0
    K = torch.zeros(S, 3, 3, dtype=torch.float32, device=dev)
    K[:, 0, 0], K[:, 1, 1], K[:, 0, 2], K[:, 1, 2], K[:, 2, 2] = cam[:, 12], cam[:, 13], cam[:, 14], cam[:, 15], 1.0
    return c2w, K


class B200GaussianSplatRenderer(torch.nn.Module):
    """`GaussianSplatRenderer` (src/models/models/rasterization.py:97-246) as far as `WorldMirror.forward` observes it:
    `render` fills `predictions["splats"]`.  The reference also rasterises every view with gsplat and then drops the
    images (rasterization.py:221-246); that dead compute is not reproduced.  Built for the configuration WorldMirror
    constructs (worldmirror.py:110-118): sh_degree 0, no offsets, residual SH, position from gs_depth + predicted camera.
    `enable_conf_filter` (off in that configuration) drops the least confident splats before the merge
    (`apply_confidence_filter`, rasterization.py:211-213, 248-299)."""

    SPLAT_KEYS = ("means", "quats", "scales", "opacities", "sh", "weights")

    def __init__(self, feature_dim: int = 256, sh_degree: int = 0, predict_offset: bool = False, predict_residual_sh: bool = True,
                 enable_prune: bool = True, voxel_size: float = 0.002, using_gtcamera_splat: bool = False,
                 render_novel_views: bool = False, enable_conf_filter: bool = False, conf_threshold_percent: float = 30.0,
                 max_gaussians: int = 5000000, debug=False, sharded: bool = False, group=None):
        super().__init__()
        if feature_dim != 256 or sh_degree != 0 or predict_offset or not predict_residual_sh:
            raise NotImplementedError("B200GaussianSplatRenderer covers WorldMirror's configuration: feature_dim 256, "
                                      "sh_degree 0, no offsets, residual SH")
        self.feature_dim, self.sh_degree, self.nums_sh = feature_dim, sh_degree, 1
        self.enable_prune, self.voxel_size = enable_prune, voxel_size
        self.enable_conf_filter, self.conf_threshold_percent, self.max_gaussians = enable_conf_filter, conf_threshold_percent, max_gaussians
        self._fws = None
        self.sharded, self.group = sharded, group  # views sharded over the ranks of `group` (None: the default group)
        # sharded: every rank merges one key range of the scene (partition_scene) instead of the whole scene W times over;
        # predictions["splats"] then holds this rank's range (WM_SPLAT_PARTITION=0 / 1 forces the replicated / partitioned merge)
        # None: automatic (more than two ranks: at two the exchange costs what the second merge does), True / False: forced
        env = os.environ.get("WM_SPLAT_PARTITION")
        self.partitioned = None if env is None else env != "0"
        self.w: Dict[str, torch.Tensor] = {}
        self._loaded = False
        self._ws = None
        from .raster import B200Rasterizer

        self.rasterizer = B200Rasterizer()  # rasterization.py:163; only used when render(..., rasterize=True)

    def prediction_cameras(self, camera_params, H: int, W: int):
        """prepare_prediction_cameras (rasterization.py:506-530): camera_params [S, 9] -> camera-to-world [S, 4, 4], K [S, 3, 3]."""
        return camera_matrices(camera_params, H, W)

    def load_state_dict(self, sd, strict: bool = True, device=None):
        """Reference keys: gs_head.0.weight [256, 128, 3, 3] (no bias), gs_head.2.weight [12, 256, 1, 1], gs_head.2.bias [12]."""
        capi.lib()
        dev = torch.device("cuda", torch.cuda.current_device()) if device is None else device
        t = sd["gs_head.0.weight"].detach().to(dev, torch.float32)
        co, ci, kh, kw = t.shape
        self.w = {"w1": t.permute(0, 2, 3, 1).reshape(co, kh * kw * ci).to(ACT).contiguous(),
                  "w2": sd["gs_head.2.weight"].detach().to(dev, torch.float32).reshape(-1, co).contiguous(),
                  "b2": sd["gs_head.2.bias"].detach().to(dev, torch.float32).contiguous()}
        assert (co, ci) == (256, 128) and self.w["w2"].shape[0] == 12
        self._loaded, self._dev = True, dev
        return torch.nn.modules.module._IncompatibleKeys([], [])

    # ------------------------------------------------------------------ pieces
    def gs_head(self, feat_nhwc: torch.Tensor) -> torch.Tensor:
        """[n, H, W, 128] fp16 channels-last -> raw Gaussian parameters [n, H, W, 12] fp32 (rasterization.py:149-153, 196-197)."""
        n, H, W, c = feat_nhwc.shape
        feat_nhwc = feat_nhwc.to(ACT).contiguous()  # no copy when it comes from B200DPTHead
        raw = torch.empty(n, H, W, 12, dtype=torch.float32, device=feat_nhwc.device)
        _k(capi.lib().wm_conv2d_project(_st(), feat_nhwc.data_ptr(), n, H, W, c, self.w["w1"].data_ptr(), None,
                                        self.w["w2"].data_ptr(), self.w["b2"].data_ptr(), 12, raw.data_ptr()))
        return raw

    def prepare_splats(self, raw, images, gs_depth, camera_params):
        """rasterization.py:389-498 for one scene: raw [n, H, W, 12], images [n, 3, H, W], gs_depth [n, H, W(, 1)],
        camera_params [n, 9] -> dict of fp32 tensors over N = n*H*W splats."""
        lib = capi.lib()
        n, H, W, _ = raw.shape
        N, dev = n * H * W, raw.device
        cam = torch.empty(n, 16, dtype=torch.float32, device=dev)
        cp = camera_params.to(dev, torch.float32).contiguous()
        _k(lib.wm_splat_cameras(_st(), cp.data_ptr(), n, H, W, cam.data_ptr()))
        out = {k: torch.empty((N, d) if d > 1 else (N,), dtype=torch.float32, device=dev)
               for k, d in (("means", 3), ("quats", 4), ("scales", 3), ("opacities", 1), ("sh", 3), ("weights", 1))}
        images, gs_depth = images.to(dev, torch.float32).contiguous(), gs_depth.contiguous()
        _k(lib.wm_splat_prepare(_st(), raw.data_ptr(), images.data_ptr(), gs_depth.data_ptr(), cam.data_ptr(), n, H, W,
                                out["means"].data_ptr(), out["quats"].data_ptr(), out["scales"].data_ptr(),
                                out["opacities"].data_ptr(), out["sh"].data_ptr(), out["weights"].data_ptr()))
        return out

    def apply_confidence_filter(self, splats, conf):
        """rasterization.py:248-299 for one scene: splats over N rows, conf with N elements (view-major, the splat order) ->
        the K = max(1, min(max_gaussians, ceil(N (100 - p) / 100))) most confident, confidences <= 1e-5 counting as -inf.
        The reference's torch.topk(sorted=False) leaves order and tie choice open; here ties go to the lower index and the
        survivors keep their original order (exact radix select, no host read: K follows from N).  Keys outside
        `mask_keys` pass through untouched, as in the reference."""
        if not self.enable_conf_filter or conf is None:
            return splats
        lib = capi.lib()
        m = splats["means"]
        N, dev = m.shape[0], m.device
        conf = conf.to(dev, torch.float32).reshape(-1).contiguous()
        if conf.numel() != N:
            raise RuntimeError(f"apply_confidence_filter: {conf.numel()} confidences for {N} splats")
        p = self.conf_threshold_percent
        keep = int(math.ceil(N * (100.0 - p) / 100.0)) if p > 0 else N
        K = max(1, min(self.max_gaussians, keep))
        need = C.c_size_t(0)
        capi.check(lib.wm_conf_filter_workspace_bytes(N, C.byref(need)))
        if self._fws is None or self._fws.numel() < need.value or self._fws.device != dev:
            self._fws = torch.empty(need.value, dtype=torch.uint8, device=dev)
        src = {k: splats[k].contiguous() for k in self.SPLAT_KEYS}
        out = {k: torch.empty((K,) + tuple(src[k].shape[1:]), dtype=torch.float32, device=dev) for k in self.SPLAT_KEYS}
        LAUNCHES[0] += 12  # init, 4 x (histogram, pick), mark, scan (CUB, counted once), gather
        capi.check(lib.wm_conf_filter_splats(_st(), N, K, conf.data_ptr(), src["means"].data_ptr(), src["quats"].data_ptr(),
                                             src["scales"].data_ptr(), src["opacities"].data_ptr(), src["sh"].data_ptr(),
                                             src["weights"].data_ptr(), self._fws.data_ptr(), self._fws.numel(),
                                             out["means"].data_ptr(), out["quats"].data_ptr(), out["scales"].data_ptr(),
                                             out["opacities"].data_ptr(), out["sh"].data_ptr(), out["weights"].data_ptr()))
        return {k: out.get(k, v) for k, v in splats.items()}

    def prune_gs(self, splats, voxel_size=None, defer: bool = False, bounds=None):
        """rasterization.py:301-387 for one scene (tensors over N splats) -> tensors over the K occupied voxels, ascending
        flat voxel index.  One host read (K) at the end, as the reference's `len(unique_voxels)`; defer=True returns the
        N-row buffers and the device-side count instead, so that the caller can keep launching work and read K later
        (`finish_prune`)."""
        lib = capi.lib()
        vs = self.voxel_size if voxel_size is None else voxel_size
        m = splats["means"]
        N, dev = m.shape[0], m.device
        out = {k: torch.empty_like(splats[k]) for k in ("means", "quats", "scales", "opacities", "sh")}
        count = torch.zeros(1, dtype=torch.int32, device=dev)
        if N == 0:  # a key range of a partitioned scene may be empty
            return (out, count) if defer else self.finish_prune(out, count)
        need = C.c_size_t(0)
        capi.check(lib.wm_prune_workspace_bytes(N, C.byref(need)))
        if self._ws is None or self._ws.numel() < need.value or self._ws.device != dev:
            self._ws = torch.empty(need.value, dtype=torch.uint8, device=dev)
        LAUNCHES[0] += 6  # bounds, keys, sort (several CUB kernels, counted once), flags, scan, merge
        args = (_st(), N, m.data_ptr(), splats["quats"].data_ptr(), splats["scales"].data_ptr(), splats["opacities"].data_ptr(),
                splats["sh"].reshape(N, 3).data_ptr(), splats["weights"].data_ptr(), float(vs), self._ws.data_ptr(),
                self._ws.numel(), out["means"].data_ptr(), out["quats"].data_ptr(), out["scales"].data_ptr(),
                out["opacities"].data_ptr(), out["sh"].data_ptr(), count.data_ptr())
        if bounds is None:
            capi.check(lib.wm_prune_splats(*args))
        else:  # the voxel-index bounds of the WHOLE scene: this call merges one key range of it
            capi.check(lib.wm_prune_splats_bounds(*args, bounds.data_ptr()))
        return (out, count) if defer else self.finish_prune(out, count)

    _PACK = (("means", 3), ("quats", 4), ("scales", 3), ("opacities", 1), ("sh", 3), ("weights", 1))

    def use_partition(self) -> bool:
        """Whether a sharded scene is merged one key range per rank (partition_scene) or whole on every rank."""
        if not self.sharded or not self.enable_prune:
            return False
        if self.partitioned is not None:
            return bool(self.partitioned)
        import torch.distributed as dist

        return dist.get_world_size(self.group) > 2

    def partition_scene(self, sp):
        """Views sharded over the ranks of `self.group`: instead of gathering all N_total splats on every rank and merging
        them W times over, every rank ends up with ONE KEY RANGE of the scene and merges only that (N_total / W splats).
          1. voxel-index bounds of the local splats, all-reduced (min / max) to the bounds of the scene;
          2. flat voxel keys of the local splats with those bounds (the keys of the single-GPU merge);
          3. W - 1 splitters = quantiles of a fixed-size sample of every rank's keys (all-gathered, so identical everywhere);
             rank r owns the keys in [splitter[r-1], splitter[r]): equal keys always meet on one rank;
          4. the local splats, stably ordered by owner, go out in ONE all-to-all of packed 15-float rows; the per-pair
             row counts are all-gathered first and read by the host (the one synchronisation of this path);
          5. what arrives is ordered by (source rank, original index) = the scene's global splat order, so the stable sort
             inside the merge accumulates every voxel in the same order as the single-GPU merge: the ranks' results,
             concatenated in rank order, equal it bit for bit (tests/sp_check.py).
        Returns (splat dict of this rank's key range, bounds)."""
        import torch.distributed as dist

        lib, group = capi.lib(), self.group
        world, rank = dist.get_world_size(group), dist.get_rank(group)
        m = sp["means"]
        n, dev = m.shape[0], m.device
        vs = float(self.voxel_size)
        bounds = torch.empty(6, dtype=torch.int64, device=dev)
        _k(lib.wm_voxel_bounds(_st(), n, m.data_ptr(), vs, bounds.data_ptr()))
        lo, hi = bounds[:3].clone(), bounds[3:].clone()
        dist.all_reduce(lo, op=dist.ReduceOp.MIN, group=group)
        dist.all_reduce(hi, op=dist.ReduceOp.MAX, group=group)
        bounds = torch.cat([lo, hi])
        keys = torch.empty(n, dtype=torch.int64, device=dev)
        _k(lib.wm_voxel_keys(_st(), n, m.data_ptr(), vs, bounds.data_ptr(), keys.data_ptr()))
        S = 4096  # sample size per rank
        samp = keys[(torch.arange(S, device=dev, dtype=torch.int64) * n) // S]
        allsamp = torch.empty(world * S, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(allsamp, samp, group=group)
        allsamp = torch.sort(allsamp).values
        splitters = allsamp[(torch.arange(1, world, device=dev, dtype=torch.int64) * (world * S)) // world].contiguous()
        dest = torch.bucketize(keys, splitters, right=True)
        order = torch.sort(dest, stable=True).indices
        counts = torch.bincount(dest, minlength=world)
        cm = torch.empty(world * world, dtype=torch.int64, device=dev)
        dist.all_gather_into_tensor(cm, counts, group=group)
        cm = cm.reshape(world, world).cpu()  # host read: cm[s, d] = rows rank s sends to rank d
        send, recv = cm[rank].tolist(), cm[:, rank].tolist()
        packed = torch.cat([sp[k].reshape(n, -1) for k, _ in self._PACK], dim=1).index_select(0, order)
        got = torch.empty(sum(recv), packed.shape[1], dtype=torch.float32, device=dev)
        dist.all_to_all_single(got, packed, recv, send, group=group)
        part, c0 = {}, 0
        for k, d in self._PACK:
            t = got[:, c0:c0 + d].contiguous()
            part[k] = t if d > 1 else t.reshape(-1)
            c0 += d
        LAUNCHES[0] += 3  # bounds, keys, (torch's sort / bucketize / gather kernels are library code)
        return part, bounds

    def gather_partitions(self, splats):
        """predictions["splats"] of a partitioned merge (this rank's key range) -> the whole merged scene on every rank,
        per-scene lists as the single-GPU renderer returns them.  One small all-gather of the row counts (host read) and
        one padded all-gather per array."""
        import torch.distributed as dist

        group = self.group
        world = dist.get_world_size(group)
        out = {k: [] for k in splats}
        for b in range(len(splats["means"])):
            K = splats["means"][b].shape[0]
            dev = splats["means"][b].device
            ks = torch.empty(world, dtype=torch.int64, device=dev)
            dist.all_gather_into_tensor(ks, torch.tensor([K], dtype=torch.int64, device=dev), group=group)
            ks = ks.cpu().tolist()
            cap = max(max(ks), 1)
            for k, v in splats.items():
                t = v[b]
                pad = torch.zeros((cap,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
                pad[:K] = t
                allp = torch.empty((world * cap,) + tuple(t.shape[1:]), dtype=t.dtype, device=dev)
                dist.all_gather_into_tensor(allp, pad, group=group)
                out[k].append(torch.cat([allp[r * cap:r * cap + ks[r]] for r in range(world)], dim=0))
        return out

    @staticmethod
    def finish_prune(out, count):
        K = int(count.item())
        res = {k: v[:K] for k, v in out.items()}
        res["sh"] = res["sh"].reshape(K, 1, 3)
        return res

    def finish(self, predictions):
        """Second half of render(..., defer=True): read the voxel counts (host sync) and slice the merged splats."""
        pending = predictions.pop("splats_pending", None)
        if pending is not None:
            predictions["splats_rows"] = [out["means"].shape[0] for out, _ in pending]  # rows of the N-row merge buffers
            merged = [self.finish_prune(out, count) for out, count in pending]
            predictions["splats"] = {k: [m[k] for m in merged] for k in ("means", "sh", "opacities", "scales", "quats")}
        return predictions

    # ------------------------------------------------------------------ reference entry point
    @torch.no_grad()
    def render(self, gs_feats, images, predictions, views=None, context_predictions=None, rasterize: bool = False,
               defer: bool = False):
        """gs_feats [B, S, 128, H, W] (as B200DPTHead(is_gsdpt=True) returns it), images [B, S, 3, H, W], predictions with
        `camera_params` [B, S, 9] and `gs_depth` [B, S, H, W, 1].  Adds predictions["splats"]: with pruning a dict of
        per-scene lists (means [K,3], sh [K,1,3], opacities [K], scales [K,3], quats [K,4]), as rasterization.py:381-386;
        without, the prepare_splats dict batched over B.  rasterize=True additionally renders every view from its predicted
        camera in chunks of four, as the reference does before it drops the images (rasterization.py:221-241), and keeps
        them as predictions["rendered_colors" | "rendered_depths" | "rendered_alphas"] [B, S, H, W, 3 | 1 | 1].
        defer=True (pruning only): no host read here; predictions["splats_pending"] holds the N-row buffers and device
        counts until `finish(predictions)` turns them into predictions["splats"]."""
        if not self._loaded:
            raise RuntimeError("B200GaussianSplatRenderer: load_state_dict() has not been called")
        if context_predictions is not None:
            raise NotImplementedError("context_predictions (training-time scale alignment) is not part of the inference path")
        raws = self.parameters_of(gs_feats, images, views)
        return self.assemble(raws, images, predictions, views, rasterize=rasterize, defer=defer)

    def parameters_of(self, gs_feats, images, views=None):
        """First half of render: the parameter head on every scene's context views -> list of raw [S, H, W, 12] (needs
        neither the cameras nor gs_depth, so `B200WorldMirror` can launch it before the camera head has finished)."""
        if not self._loaded:
            raise RuntimeError("B200GaussianSplatRenderer: load_state_dict() has not been called")
        B, S_all = images.shape[:2]
        S = int(views["context_nums"]) if views is not None and "context_nums" in views else S_all
        feat = gs_feats.permute(0, 1, 3, 4, 2)  # back to channels-last: a view when it came from B200DPTHead
        return [self.gs_head(feat[b, :S]) for b in range(B)]

    def assemble(self, raws, images, predictions, views=None, rasterize: bool = False, defer: bool = False):
        """Second half of render: prepare_splats, the gather across ranks when views are sharded, the voxel merge."""
        B = images.shape[0]
        S = raws[0].shape[0]
        scenes = []
        for b in range(B):
            raw = raws[b]
            sp = self.prepare_splats(raw, images[b, :S], predictions["gs_depth"][b, :S], predictions["camera_params"][b, :S])
            conf_on = self.enable_conf_filter and "depth_conf" in predictions
            if self.use_partition() and not rasterize and not conf_on:
                # every rank keeps and merges one key range of the scene (partition_scene)
                part, bounds = self.partition_scene(sp)
                scenes.append((part, bounds))
                continue
            if self.sharded:  # views are sharded over ranks: every rank needs all splats of the scene
                import torch.distributed as dist

                world = dist.get_world_size(self.group)
                for k in self.SPLAT_KEYS:
                    parts = [torch.empty_like(sp[k]) for _ in range(world)]
                    dist.all_gather(parts, sp[k], group=self.group)
                    sp[k] = torch.cat(parts, dim=0)
            sp["gs_feats"] = raw.reshape(-1, 12)
            if self.enable_conf_filter and "depth_conf" in predictions:  # rasterization.py:211-213 (the depth head's confidence)
                conf = predictions["depth_conf"][b, :S].to(torch.float32).contiguous()
                if self.sharded:
                    parts = [torch.empty_like(conf) for _ in range(world)]
                    dist.all_gather(parts, conf, group=self.group)
                    conf = torch.cat(parts, dim=0)
                sp = self.apply_confidence_filter(sp, conf)
            scenes.append(sp)
        if scenes and isinstance(scenes[0], tuple):  # partitioned: (this rank's key range, the scene's bounds)
            if defer:
                predictions["splats_pending"] = [self.prune_gs(sp, defer=True, bounds=bd) for sp, bd in scenes]
            else:
                merged = [self.prune_gs(sp, bounds=bd) for sp, bd in scenes]
                predictions["splats"] = {k: [m[k] for m in merged] for k in ("means", "sh", "opacities", "scales", "quats")}
            predictions["splats_partitioned"] = True  # this rank's key range; gather_partitions() assembles the scene
        elif self.enable_prune and defer and not rasterize:
            predictions["splats_pending"] = [self.prune_gs(sp, defer=True) for sp in scenes]
        elif self.enable_prune:
            merged = [self.prune_gs(sp) for sp in scenes]
            predictions["splats"] = {k: [m[k] for m in merged] for k in ("means", "sh", "opacities", "scales", "quats")}
            if rasterize:  # views sharded over ranks: every rank holds the whole merged scene and renders its own views
                H, W = images.shape[-2:]
                outs = [[], [], []]
                for b, m in enumerate(merged):
                    c2w, K = self.prediction_cameras(predictions["camera_params"][b], H, W)
                    parts = [self.rasterizer.rasterize_splats(m["means"], m["quats"], m["scales"], m["opacities"], m["sh"], c2w[i:i + 4],
                                                              K[i:i + 4], W, H, sh_degree=0) for i in range(0, c2w.shape[0], 4)]
                    for k in range(3):
                        outs[k].append(torch.cat([p[k] for p in parts], dim=0))
                for k, name in enumerate(("rendered_colors", "rendered_depths", "rendered_alphas")):
                    predictions[name] = torch.stack(outs[k], dim=0)
        else:
            spl = {k: torch.stack([sp[k] for sp in scenes], dim=0) for k in self.SPLAT_KEYS + ("gs_feats",)}
            spl["sh"] = spl["sh"].reshape(B, -1, 1, 3)
            spl["residual_sh"] = spl["gs_feats"][..., 8:11].reshape(B, -1, 1, 3)
            predictions["splats"] = spl
        return predictions


class B200WorldMirror(torch.nn.Module):
    """`WorldMirror.forward` (src/models/models/worldmirror.py:120-196) on the B200 path: aggregator, camera head and the
    depth / points / normals DPT heads, same `views` dictionary in, same prediction keys out (`camera_params`,
    `camera_poses` [B,S,4,4] camera-to-world, `camera_intrs` [B,S,3,3], `depth`, `depth_conf`, `pts3d`, `pts3d_conf`,
    `normals`, `normals_conf`; the key set is pinned against the reference's own forward, tests/golden/c1_preds.npz).  With any cond_flag set the priors are extracted
    from `views` (`camera_pose`, `depthmap`, `camera_intrinsics`) by `priors.extract_priors`, as `WorldMirror.forward`
    does (worldmirror.py:133-143); `priors` may instead pass the already normalised (depth, rays, poses).  With the
    Gaussian branch (`gs_head`, `gs_renderer`; worldmirror.py:202-217) the predictions also carry `gs_depth`,
    `gs_depth_conf` and `splats`."""

    def __init__(self, aggregator, cam_head=None, depth_head=None, pts_head=None, norm_head=None, gs_head=None, gs_renderer=None):
        super().__init__()
        self.visual_geometry_transformer = aggregator
        self.cam_head, self.depth_head, self.pts_head, self.norm_head = cam_head, depth_head, pts_head, norm_head
        self.gs_head, self.gs_renderer = gs_head, gs_renderer
        self.enable_gs = gs_head is not None and gs_renderer is not None
        self._side = None
        self.share_ln = os.environ.get("WM_SHARE_LN", "1") != "0"  # one LayerNorm(2048) of the token rows for all DPT heads
        self.head_streams = os.environ.get("WM_HEAD_STREAMS", "0") == "1"
        self._hstreams = []
        self.check_fp16_range = True  # raise Fp16RangeError when a DPT head's fp16 map saturates
        # views sharded over ranks: camera head on the local views (WM_CAMERA_SHARDED=0: all views on every rank)
        self.camera_sharded = os.environ.get("WM_CAMERA_SHARDED", "1") != "0"
        self.splat_readback = None  # (first row, rows): the slice of the merged splat buffers this rank copies to host_out

    def _camera(self, token_list):
        agg = self.visual_geometry_transformer
        if getattr(agg, "world", 1) > 1:
            if self.camera_sharded:
                # every layer but the trunk attention is per view: each rank runs its own views and the q|k|v rows of
                # the trunk blocks are all-gathered (B200CameraHead._run)
                return self.cam_head(token_list, shard=(agg._sp_group, agg.world, agg.rank))[-1]
            # replicated: gather every rank's camera tokens [S_local, 2048] and run all views on every rank
            import torch.distributed as dist

            mine = token_list[-1][:, :, :1].contiguous()
            parts = [torch.empty_like(mine) for _ in range(agg.world)]
            dist.all_gather(parts, mine, group=agg._sp_group)
            s_loc = mine.shape[1]
            cam = self.cam_head([torch.cat(parts, dim=1)])[-1]
            return cam[:, agg.rank * s_loc:(agg.rank + 1) * s_loc].contiguous()
        return self.cam_head(token_list)[-1]

    @classmethod
    def from_state_dicts(cls, aggregator_sd, heads_sd, aggregator=None, **agg_kwargs):
        from .aggregator import B200VisualGeometryTransformer

        agg = aggregator
        if agg is None:
            agg = B200VisualGeometryTransformer(**agg_kwargs)
            agg.load_state_dict(aggregator_sd)

        def sub(p):
            return {k[len(p):]: v for k, v in heads_sd.items() if k.startswith(p)}

        cam = B200CameraHead(dim_in=2048)
        cam.load_state_dict(sub("cam_head."))
        heads = {}
        for name, od, act in (("depth_head", 2, "exp+expp1"), ("pts_head", 4, "inv_log+expp1"), ("norm_head", 4, "norm+expp1")):
            h = B200DPTHead(dim_in=2048, output_dim=od, activation=act)
            h.load_state_dict(sub(name + "."))
            heads[name] = h
        gs_head = gs_renderer = None
        if any(k.startswith("gs_renderer.") for k in heads_sd):  # enable_gs (worldmirror.py:101-118)
            gs_head = B200DPTHead(dim_in=2048, output_dim=2, activation="exp+expp1", features=256, is_gsdpt=True)
            gs_head.load_state_dict(sub("gs_head."))
            gs_renderer = B200GaussianSplatRenderer(sh_degree=0, predict_offset=False, predict_residual_sh=True, enable_prune=True,
                                                    voxel_size=0.002, using_gtcamera_splat=True, render_novel_views=True,
                                                    sharded=getattr(agg, "world", 1) > 1, group=getattr(agg, "_sp_group", None))
            gs_renderer.load_state_dict(sub("gs_renderer."))
        return cls(agg, cam, heads["depth_head"], heads["pts_head"], heads["norm_head"], gs_head, gs_renderer)

    @torch.no_grad()
    def forward(self, views, cond_flags=(0, 0, 0), priors=None, host_out=None):
        """host_out: optional dict of PINNED host tensors keyed by prediction name (and "splats.means" | "splats.sh" |
        "splats.opacities" | "splats.scales" | "splats.quats", sized for views * H * W rows of one scene): every prediction is
        copied to its buffer on a copy stream as soon as the head that produces it has finished, underneath the heads
        that are still running; the call returns after enqueueing, the caller synchronises the current stream."""
        imgs = views["img"]
        if sum(cond_flags) > 0:
            if priors is None:  # worldmirror.py:136-138: priors come from the views dictionary
                from .priors import extract_priors

                agg = self.visual_geometry_transformer
                priors = extract_priors(views, group=getattr(agg, "_sp_group", None), world=getattr(agg, "world", 1),
                                        rank=getattr(agg, "rank", 0))
            with capi.nvtx_range("wm.aggregator"):
                token_list, psi = self.visual_geometry_transformer(imgs, priors, cond_flags=list(cond_flags))
        else:
            with capi.nvtx_range("wm.aggregator"):
                token_list, psi = self.visual_geometry_transformer(imgs)
        preds = {}
        sat_now = False
        main = torch.cuda.current_stream(imgs.device)
        dpt = [h for h in (self.gs_head, self.depth_head, self.pts_head, self.norm_head) if h is not None]
        ln_cache = {} if (self.share_ln and not self.head_streams and len({h.dim_in for h in dpt}) == 1) else None
        for h in dpt:
            h.ln_cache = ln_cache  # the normalised token rows are the same for every head (same stream: keyed by it)
        copy = None
        if host_out is not None:
            if getattr(self, "_copy", None) is None or self._copy.device != imgs.device:
                self._copy = torch.cuda.Stream(device=imgs.device)
            copy = self._copy

        def emit(name, t, src=None):
            if copy is None or name not in host_out:
                return
            if name.startswith("splats.") and self.splat_readback is not None:
                # the merged scene is identical on every rank: each one reads back its share of the rows
                r0, nr = self.splat_readback
                t = t[r0:r0 + nr]
            copy.wait_stream(main if src is None else src)
            with torch.cuda.stream(copy):
                dst = host_out[name]
                ne = min(t.numel(), dst.numel())  # (a key range of a partitioned scene may exceed the caller's estimate)
                dst.view(-1)[:ne].copy_(t.reshape(-1)[:ne], non_blocking=True)
            t.record_stream(copy)

        side = None
        if self.cam_head is not None:
            # The camera head is a chain of ~137 tiny kernels (latency-bound), the DPT heads are throughput-bound: run the
            # camera head on a high-priority side stream underneath them.
            if self._side is None or self._side.device != imgs.device:
                self._side = torch.cuda.Stream(device=imgs.device, priority=-1)
            side = self._side
            side.wait_stream(main)
            with torch.cuda.stream(side), capi.nvtx_range("wm.camera_head"):
                preds["camera_params"] = self._camera(token_list)
                # worldmirror.py:161-175: camera-to-world poses and intrinsics of the predicted cameras
                mats = [camera_matrices(cp, imgs.shape[-2], imgs.shape[-1]) for cp in preds["camera_params"]]
                preds["camera_poses"] = torch.stack([m[0] for m in mats], dim=0)
                preds["camera_intrs"] = torch.stack([m[1] for m in mats], dim=0)
        if self.enable_gs:
            # The Gaussian branch goes first.  Its DPT head and the parameter head need no camera: they run on the main
            # stream.  Everything that does (prepare_splats, the gather across ranks, the voxel merge) continues on the side
            # stream behind the camera head, so the main stream never waits for the camera, and the splats - the largest
            # result - are merged and read back underneath the three remaining DPT heads.
            if self.cam_head is None:
                raise RuntimeError("the Gaussian branch positions its splats with the predicted cameras: cam_head is required")
            with capi.nvtx_range("wm.gs_head"):
                gs_feat, preds["gs_depth"], preds["gs_depth_conf"] = self.gs_head(token_list, images=imgs, patch_start_idx=psi)
            emit("gs_depth", preds["gs_depth"])
            emit("gs_depth_conf", preds["gs_depth_conf"])
            raws = self.gs_renderer.parameters_of(gs_feat, imgs, views)
            del gs_feat
            if self.gs_renderer.enable_conf_filter and self.depth_head is not None:
                # the confidence filter reads the depth head's confidence (rasterization.py:211-213): that head goes first
                preds["depth"], preds["depth_conf"] = self.depth_head(token_list, images=imgs, patch_start_idx=psi)
                emit("depth", preds["depth"])
                emit("depth_conf", preds["depth_conf"])
            gs_ready = main.record_event()

            def splat_chain(preds):
                side.wait_event(gs_ready)
                with torch.cuda.stream(side):
                    for t in raws + [preds["gs_depth"], imgs] + ([preds["depth_conf"]] if "depth_conf" in preds else []):
                        t.record_stream(side)
                    emit("camera_params", preds["camera_params"], src=side)
                    with capi.nvtx_range("wm.splats"):
                        preds = self.gs_renderer.assemble(raws, imgs, preds, views, defer=True)
                    if "splats_pending" in preds and len(preds["splats_pending"]) == 1:
                        for k, t in preds["splats_pending"][0][0].items():
                            emit("splats." + k, t, src=side)
                return preds

            # The partitioned merge of a sharded scene reads one small count matrix on the host in the middle of the chain:
            # the remaining heads are enqueued first, so the main stream has work while the host waits for the side stream.
            late_chain = self.gs_renderer.use_partition()
            if not late_chain:
                preds = splat_chain(preds)
        used = []
        # The depth head goes LAST: whatever the last head produces is copied to the host with nothing left to hide the copy
        # behind, and depth + confidence are half the bytes of points or normals (17 against 34 MB for 8 views of 518 x 518).
        for i, (name, head) in enumerate((("pts3d", self.pts_head), ("normals", self.norm_head), ("depth", self.depth_head))):
            if head is not None and name not in preds:
                if self.head_streams:
                    # the three heads are independent and their low-resolution stages launch fewer CTAs than there are
                    # SMs: one stream per head lets the GPU overlap them (WM_HEAD_STREAMS=1)
                    if len(self._hstreams) <= i or self._hstreams[i].device != imgs.device:
                        self._hstreams = [torch.cuda.Stream(device=imgs.device) for _ in range(3)]
                    hs = self._hstreams[i]
                    hs.wait_stream(main)
                    with torch.cuda.stream(hs):
                        preds[name], preds[name + "_conf"] = head(token_list, images=imgs, patch_start_idx=psi)
                    used.append((hs, name))
                    emit(name, preds[name], src=hs)
                    emit(name + "_conf", preds[name + "_conf"], src=hs)
                else:
                    with capi.nvtx_range("wm.dpt_" + name):
                        preds[name], preds[name + "_conf"] = head(token_list, images=imgs, patch_start_idx=psi)
                    emit(name, preds[name])
                    emit(name + "_conf", preds[name + "_conf"])
        for hs, name in used:
            main.wait_stream(hs)
            preds[name].record_stream(main)
            preds[name + "_conf"].record_stream(main)
        if self.enable_gs:
            if late_chain:
                preds = splat_chain(preds)
            del raws
        if side is not None:
            main.wait_stream(side)
            for k in ("camera_params", "camera_poses", "camera_intrs"):
                preds[k].record_stream(main)
            if not self.enable_gs:
                emit("camera_params", preds["camera_params"])
        if self.enable_gs:
            for out, count in preds.get("splats_pending", []):
                for t in list(out.values()) + [count]:
                    t.record_stream(main)
            preds = self.gs_renderer.finish(preds)  # host read of the voxel count: the stream is drained here
            sat_now = True  # ... so the saturation count copied below is valid right away
            if hasattr(self.visual_geometry_transformer, "raise_on_peer_timeout"):
                self.visual_geometry_transformer.raise_on_peer_timeout()
        for h in dpt:
            h.ln_cache = None
        if copy is not None:
            main.wait_stream(copy)
        if self.check_fp16_range and any(h is not None for h in (self.depth_head, self.pts_head, self.norm_head, self.gs_head)):
            # fp16 maps in the DPT heads: a forward whose activations left the fp16 range must not pass silently.  With the
            # Gaussian branch the stream has just been drained, so this forward's count is read now; otherwise the copy
            # is only enqueued and the PREVIOUS forward's count is checked (no extra synchronisation on the hot path;
            # `fp16_saturation()` checks explicitly).
            n_sat = fp16_saturation(imgs.device, reset=True, sync=sat_now)
            if n_sat:
                raise Fp16RangeError(f"{n_sat} (pixel, 32-channel) entries of the DPT heads' fp16 maps left the fp16 range "
                                     f"in {'this' if sat_now else 'the previous'} forward: the predictions are not valid")
        # the order in which the reference fills its dictionary (worldmirror.py:161-217), whatever order the heads ran in
        first = ("camera_params", "camera_poses", "camera_intrs", "depth", "depth_conf", "pts3d", "pts3d_conf", "normals", "normals_conf")
        preds = {**{k: preds[k] for k in first if k in preds}, **{k: v for k, v in preds.items() if k not in first}}
        return preds
