"""Drop-in for the reference aggregator module, `VisualGeometryTransformer`
(src/models/models/visual_transformer.py:20-416 of the reference), backed by libwm_b200.

Same constructor keywords, same `forward(images, priors, cond_flags)` signature, same return value
`(list of 4 tensors [B, S, P, 2048] fp32, patch_start_idx=7)`, same `state_dict` keys on the way in, same
error behaviour (`ValueError` for channels != 3, visual_transformer.py:272-273; assertion for H, W not
multiples of 14, patch_embed.py:67-68).  Assign an instance to `WorldMirror.visual_geometry_transformer`
(worldmirror.py:50-61) and the rest of the reference model runs unchanged.

PyTorch is plumbing here: device memory, the current stream and (for sequence parallelism) the process
group used to hand out the NCCL id.  All arithmetic happens in the CUDA library; there is no CPU path.
"""
from __future__ import annotations

import ctypes as C
import math
from typing import Dict, List, Optional, Sequence, Tuple

import torch
import torch.nn.functional as F

from . import capi
from .weights import state_dict_schema

PATCH = 14
EMBED = 1024
PATCH_START = 7
DINO_GRID = 37


def _ptr(t: Optional[torch.Tensor]) -> Optional[int]:
    return None if t is None else t.data_ptr()


class B200VisualGeometryTransformer(torch.nn.Module):
    """B200-native `VisualGeometryTransformer`.  Only the published WorldMirror configuration is built
    (ViT-L/14 DINOv2 with 4 registers, C=1024, 16 heads, QK-norm, RoPE base 100, enable_cond=True);
    other keyword values raise instead of silently computing something else."""

    def __init__(self, img_size=518, patch_size=14, embed_dim=1024, depth=24, num_heads=16, mlp_ratio=4.0,
                 num_register_tokens=4, block_fn=None, qkv_bias=True, proj_bias=True, ffn_bias=True,
                 patch_embed="dinov2_vitl14_reg", qk_norm=True, rope_freq=100, init_values=0.01, enable_cond=True,
                 sampling_strategy="uniform", fixed_patch_embed=False, condition_strategy=("token", "pow3r", "token"),
                 intermediate_idxs: Sequence[int] = (4, 11, 17, 23), dino_depth: int = 24, **_ignored):
        super().__init__()
        fixed = dict(patch_size=(patch_size, 14), embed_dim=(embed_dim, 1024), num_heads=(num_heads, 16),
                     mlp_ratio=(float(mlp_ratio), 4.0), num_register_tokens=(num_register_tokens, 4),
                     qkv_bias=(qkv_bias, True), proj_bias=(proj_bias, True), ffn_bias=(ffn_bias, True),
                     patch_embed=(patch_embed, "dinov2_vitl14_reg"), qk_norm=(qk_norm, True), rope_freq=(rope_freq, 100),
                     enable_cond=(enable_cond, True))
        for name, (got, want) in fixed.items():
            if got != want:
                raise NotImplementedError(f"libwm_b200 implements {name}={want!r} only (got {got!r})")
        if list(condition_strategy) != ["token", "pow3r", "token"]:
            raise NotImplementedError("libwm_b200 implements condition_strategy=['token','pow3r','token'] only")
        idxs = [int(i) for i in intermediate_idxs if int(i) < depth]
        if len(idxs) > 4 or idxs != sorted(set(idxs)):
            raise ValueError("intermediate_idxs must be at most 4 increasing layer indices")
        self.depth, self.dino_depth, self.intermediate_idxs = int(depth), int(dino_depth), idxs
        self.enable_cond = True
        self.patch_size = PATCH
        self.patch_start_idx = PATCH_START
        self._handle = None
        self._loaded = False
        self._pos_embed = None  # fp32 copy of patch_embed.pos_embed for non-native grids
        self._pos_grid = (DINO_GRID, DINO_GRID)
        self._workspace: Dict[Tuple[int, int, int], torch.Tensor] = {}
        self.rank, self.world = 0, 1
        self._sp_group, self._p2p_rows = None, 0

    # ------------------------------------------------------------------ library handle
    def _ensure_handle(self):
        if self._handle is None:
            lib = capi.lib()
            cfg = capi.WmConfig()
            cfg.depth, cfg.dino_depth, cfg.n_intermediate = self.depth, self.dino_depth, len(self.intermediate_idxs)
            for i, v in enumerate(self.intermediate_idxs):
                cfg.intermediate[i] = v
            h = C.c_void_p()
            capi.check(lib.wm_create(C.byref(cfg), C.byref(h)))
            self._handle = h
        return self._handle

    def __del__(self):
        try:
            if getattr(self, "_handle", None) is not None:
                capi.lib().wm_destroy(self._handle)
                self._handle = None
        except Exception:
            pass

    # ------------------------------------------------------------------ weights
    def load_state_dict(self, state_dict, strict: bool = True, assign: bool = False):  # noqa: D401
        """Takes the reference `state_dict` unchanged: same keys, same shapes."""
        schema = state_dict_schema(self.depth, self.dino_depth)
        missing = [k for k in schema if k not in state_dict]
        unexpected = [k for k in state_dict if k not in schema]
        if missing or (strict and unexpected and not _only_deeper_blocks(unexpected)):
            raise RuntimeError(f"Error(s) in loading state_dict: missing keys {missing[:5]} ({len(missing)} total), "
                               f"unexpected keys {unexpected[:5]} ({len(unexpected)} total)")
        lib, h = capi.lib(), self._ensure_handle()
        for key, shape in schema.items():
            t = state_dict[key]
            if tuple(t.shape) != tuple(shape):
                raise RuntimeError(f"size mismatch for {key}: expected {tuple(shape)}, got {tuple(t.shape)}")
            if key.endswith("mask_token"):
                continue  # unused at inference (vision_transformer.py:213-214)
            t = t.detach().to(torch.float32).contiguous()
            dims = (C.c_int64 * t.dim())(*t.shape)
            capi.check(lib.wm_load_weight(h, key.encode(), t.data_ptr(), dims, t.dim()))
        capi.check(lib.wm_finalize_weights(h))
        self._pos_embed = state_dict["patch_embed.pos_embed"].detach().to(torch.float32).cpu().clone()
        self._pos_grid = (DINO_GRID, DINO_GRID)
        self._loaded = True
        return torch.nn.modules.module._IncompatibleKeys([], unexpected)

    def _ensure_pos_table(self, gh: int, gw: int):
        """interpolate_pos_encoding (vision_transformer.py:175-207): bicubic + antialias resample of the 37x37
        table, fp32, done once per grid and handed to the library."""
        if (gh, gw) == self._pos_grid:
            return
        pe = self._pos_embed[0]
        if gh == DINO_GRID and gw == DINO_GRID:
            table = pe
        else:
            m = int(math.isqrt(pe.shape[0] - 1))
            grid = pe[1:].reshape(1, m, m, EMBED).permute(0, 3, 1, 2)
            grid = F.interpolate(grid, size=(gh, gw), mode="bicubic", antialias=True)
            table = torch.cat([pe[:1], grid.permute(0, 2, 3, 1).reshape(gh * gw, EMBED)], dim=0)
        table = table.contiguous()
        capi.check(capi.lib().wm_set_dino_pos_table(self._handle, gh, gw, table.data_ptr()))
        self._pos_grid = (gh, gw)

    # ------------------------------------------------------------------ sequence parallelism
    def init_sequence_parallel(self, group=None):
        """One process per GPU: frame attention stays local, global attention all-gathers K|V through NCCL
        (SURVEY.md section 8e).  The NCCL id travels over the existing torch.distributed group."""
        import torch.distributed as dist

        rank, world = dist.get_rank(group), dist.get_world_size(group)
        if world == 1:
            return
        from .parallel import broadcast_unique_id

        path = nccl_library_path()
        uid = None
        if rank == 0:
            buf = (C.c_uint8 * 128)()
            capi.check(capi.lib().wm_nccl_unique_id(path.encode(), buf))
            uid = bytes(buf)
        dev = torch.device("cuda", torch.cuda.current_device()) if dist.get_backend(group) == "nccl" else None
        raw = (C.c_uint8 * 128)(*broadcast_unique_id(uid, group, dev))
        capi.check(capi.lib().wm_nccl_init(self._ensure_handle(), path.encode(), raw, rank, world))
        self.rank, self.world = rank, world
        self._sp_group = group
        self._p2p_rows = 0

    def _ensure_peer_gather(self, rows_per_rank: int, device):
        """Peer-memory K|V exchange (include/wm_b200.h, wm_p2p_*): every rank allocates its gather block, the
        cudaIpcMemHandle_t's travel over the torch.distributed group, every rank maps its peers' blocks.  Collective:
        all ranks reach this with the same `rows_per_rank` (they run the same forwards).  WM_SP_MODE=nccl keeps
        the plain ncclAllGather path."""
        import os

        import torch.distributed as dist

        if os.environ.get("WM_SP_MODE", "p2p") == "nccl" or rows_per_rank <= self._p2p_rows:
            return
        group = self._sp_group
        torch.cuda.synchronize(device)
        dist.barrier(group)  # nobody is still pushing into a block that is about to be freed
        mine = (C.c_uint8 * 64)()
        capi.check(capi.lib().wm_p2p_alloc(self._handle, self.rank, self.world, rows_per_rank, mine))
        t = torch.tensor(list(bytes(mine)), dtype=torch.uint8)
        if dist.get_backend(group) == "nccl":
            t = t.to(device)
        parts = [torch.empty_like(t) for _ in range(self.world)]
        dist.all_gather(parts, t, group=group)
        raw = b"".join(bytes(x.cpu().tolist()) for x in parts)
        capi.check(capi.lib().wm_p2p_attach(self._handle, (C.c_uint8 * len(raw)).from_buffer_copy(raw)))
        dist.barrier(group)
        self._p2p_rows = rows_per_rank

    def peer_gather_status(self):
        """(enabled, timed_out) of the peer-memory exchange; timed_out means a forward ran on stale keys.  Synchronous
        (one 4-byte device read) and clears the condition."""
        en, to = C.c_int(), C.c_int()
        capi.check(capi.lib().wm_p2p_status(self._handle, C.byref(en), C.byref(to)))
        return bool(en.value), bool(to.value)

    def raise_on_peer_timeout(self):
        """Call where the host synchronises anyway (B200WorldMirror.forward does): raises if a K|V shard of a peer did not
        arrive in time during the forwards since the last check.  Without this call the library itself reports the
        condition from the NEXT forward (wm_forward returns WM_ERR_STATE once), so it can never go unnoticed."""
        if self.world > 1:
            enabled, timed_out = self.peer_gather_status()
            if enabled and timed_out:
                raise RuntimeError("sequence-parallel K|V exchange timed out: a peer's shard did not arrive within 4 s and the "
                                   "forward finished on stale keys (a rank is slow, desynchronised or has crashed)")

    # ------------------------------------------------------------------ forward
    def _workspace_for(self, n: int, H: int, W: int, device) -> torch.Tensor:
        key = (n, H, W)
        ws = self._workspace.get(key)
        if ws is None or ws.device != device:
            nbytes = capi.lib().wm_workspace_bytes(self._handle, n, H, W, self.world)
            if nbytes == 0:
                raise ValueError(f"invalid geometry n_views={n} H={H} W={W}")
            # one buffer, as large as the largest shape seen so far: a smaller shape carves its workspace out of it, so
            # callers that alternate between shapes do not re-allocate, and peak memory stays bounded by the largest one
            big = next((t for t in self._workspace.values() if t.device == device and t.numel() >= nbytes), None)
            if big is None:
                self._workspace.clear()
                big = torch.empty(nbytes, dtype=torch.uint8, device=device)
            ws = big
            self._workspace[key] = ws
        return ws

    @torch.no_grad()
    def forward(self, images: torch.Tensor, priors=None, cond_flags: Sequence[int] = (0, 0, 0), ctx_frames=None):
        """images [B, S, 3, H, W] in [0, 1]; priors = (depth [B,S,H,W], rays [B,S,4], poses [B,S,7]) or None;
        cond_flags = [pose, depth, rays].  Under sequence parallelism S is this rank's share of the views.
        `ctx_frames` is accepted and ignored exactly like the reference."""
        if images.dim() != 5:
            raise ValueError(f"Expected images of shape [B, S, 3, H, W], got {tuple(images.shape)}")
        b, s, ch, H, W = images.shape
        if ch != 3:
            raise ValueError(f"Expected 3 input channels, got {ch}")
        assert H % PATCH == 0, f"Input image height {H} is not a multiple of patch height {PATCH}"
        assert W % PATCH == 0, f"Input image width {W} is not a multiple of patch width: {PATCH}"
        if not self._loaded:
            raise RuntimeError("B200VisualGeometryTransformer: load_state_dict() has not been called")
        if not images.is_cuda:
            raise RuntimeError("B200VisualGeometryTransformer needs CUDA tensors (there is no CPU path)")
        device = images.device
        depth_maps, ray_dirs, poses = priors if priors is not None else (None, None, None)
        flags = [int(f) for f in cond_flags]
        gh, gw = H // PATCH, W // PATCH
        P = PATCH_START + gh * gw
        lib, h = capi.lib(), self._handle
        outs = [torch.empty(b, s, P, 2 * EMBED, dtype=torch.float32, device=device) for _ in self.intermediate_idxs]
        with torch.cuda.device(device):
            self._ensure_pos_table(gh, gw)
            ws = self._workspace_for(s, H, W, device)
            if self.world > 1:
                self._ensure_peer_gather(s * P, device)
            stream = torch.cuda.current_stream(device).cuda_stream
            cflags = (C.c_int * 3)(*flags)
            for bi in range(b):
                img = images[bi].to(torch.float32).contiguous()
                d = None if depth_maps is None else depth_maps[bi].reshape(s, H, W).to(device, torch.float32).contiguous()
                r = None if ray_dirs is None else ray_dirs[bi].reshape(s, 4).to(device, torch.float32).contiguous()
                p = None if poses is None else poses[bi].reshape(s, 7).to(device, torch.float32).contiguous()
                optrs = (C.c_void_p * 4)(*[o[bi].data_ptr() for o in outs] + [None] * (4 - len(outs)))
                capi.check(lib.wm_forward(h, stream, img.data_ptr(), _ptr(d), _ptr(r), _ptr(p), cflags, s, H, W,
                                          self.rank * s, optrs, ws.data_ptr(), ws.numel()))
        return outs, PATCH_START

    @torch.no_grad()
    def forward_tokens(self, tokens: torch.Tensor, gh: int, gw: int):
        """The alternating-attention stack alone (visual_transformer.py:309-339) on assembled tokens
        [S, P, 1024] fp32 (BASELINE.json config 5).  Returns the list of [1, S, P, 2048] outputs."""
        s, P, c = tokens.shape
        if c != EMBED or P != PATCH_START + gh * gw:
            raise ValueError(f"tokens must be [S, {PATCH_START + gh * gw}, {EMBED}], got {tuple(tokens.shape)}")
        if not self._loaded:
            raise RuntimeError("B200VisualGeometryTransformer: load_state_dict() has not been called")
        device = tokens.device
        x = tokens.to(torch.float32).contiguous().clone()
        outs = [torch.empty(1, s, P, 2 * EMBED, dtype=torch.float32, device=device) for _ in self.intermediate_idxs]
        with torch.cuda.device(device):
            ws = self._workspace_for(s, gh * PATCH, gw * PATCH, device)
            if self.world > 1:
                self._ensure_peer_gather(s * P, device)
            optrs = (C.c_void_p * 4)(*[o.data_ptr() for o in outs] + [None] * (4 - len(outs)))
            capi.check(capi.lib().wm_forward_tokens(self._handle, torch.cuda.current_stream(device).cuda_stream,
                                                    x.data_ptr(), s, gh, gw, optrs, ws.data_ptr(), ws.numel()))
        return outs


def _only_deeper_blocks(keys) -> bool:
    """A truncated stack (depth < 24) may be fed the full reference state_dict: extra block keys are fine."""
    return all(k.startswith(("frame_blocks.", "global_blocks.", "patch_embed.blocks.")) for k in keys)


def nccl_library_path() -> str:
    """The libnccl.so.2 torch itself uses (nvidia-nccl wheel), else the system one."""
    import glob
    import os

    import torch as _t

    site = os.path.dirname(os.path.dirname(_t.__file__))
    for pat in (os.path.join(site, "nvidia", "nccl", "lib", "libnccl.so.2"), "/usr/lib/x86_64-linux-gnu/libnccl.so.2"):
        hits = glob.glob(pat)
        if hits:
            return hits[0]
    raise RuntimeError("libnccl.so.2 not found")


def install_into(world_mirror, state_dict=None):
    """Swap the B200 aggregator underneath a reference `WorldMirror` (worldmirror.py:50-61): heads, priors
    extraction and the public `forward(views, cond_flags)` stay byte-identical."""
    ref = world_mirror.visual_geometry_transformer
    agg = B200VisualGeometryTransformer(depth=len(ref.frame_blocks), intermediate_idxs=ref.intermediate_idxs)
    agg.load_state_dict(state_dict if state_dict is not None else ref.state_dict())
    world_mirror.visual_geometry_transformer = agg
    return agg
