#!/usr/bin/env python
"""Same-box comparator: what stock PyTorch (cuBLAS GEMMs, its SDPA backends) does on this B200 for the shapes of the
hot path -- the reference ships no custom kernel for the path, so "PyTorch's own SDPA / cuBLAS
on the same box" is the backend it would run on.  Nothing here is product code and nothing here is timed as ours.

  kernel level : F.scaled_dot_product_attention bf16 on [1,16,11008,64] (global attention, 8 views) and
                 [8,16,1376,64] (frame attention); F.linear on the four GEMM shapes of a block at M = 11008.
  model level  : the operator sequence of the reference aggregator (vision_transformer.py:253-259 DINOv2 blocks,
                 visual_transformer.py:309-339 alternating frame / global blocks; block.py:91-92, attention.py:48-69,
                 rope.py:148-181, mlp.py:29-35) built from nn.Linear / nn.LayerNorm / F.scaled_dot_product_attention /
                 nn.GELU with random weights, under torch.autocast('cuda', bf16) exactly as infer.py:157-159 runs the
                 model.  Shapes and op count match the reference; values are irrelevant for timing.

    python tools/bench_torch_gpu.py [--views 8] [--iters 5] [--out gpurun_out/torch_gpu_baseline.json]
"""
import argparse
import json
import sys

import torch
import torch.nn as nn
import torch.nn.functional as F

C, HEADS, HD, MLP = 1024, 16, 64, 4096


def time_ms(fn, iters, warmup=3):
    for _ in range(warmup):
        fn()
    torch.cuda.synchronize()
    a, b = torch.cuda.Event(enable_timing=True), torch.cuda.Event(enable_timing=True)
    a.record()
    for _ in range(iters):
        fn()
    b.record()
    torch.cuda.synchronize()
    return a.elapsed_time(b) / iters


class Rope2D(nn.Module):
    """Same op sequence as rope.py:148-181 (chunk, two embedding lookups per axis, rotate-half via cat, mul/add)."""

    def __init__(self, base=100.0):
        super().__init__()
        self.base, self.cache = base, {}

    def tables(self, dim, n, device, dtype):
        key = (dim, n, device, dtype)
        if key not in self.cache:
            inv = 1.0 / (self.base ** (torch.arange(0, dim, 2, device=device).float() / dim))
            ang = torch.einsum("i,j->ij", torch.arange(n, device=device, dtype=inv.dtype), inv).to(dtype)
            ang = torch.cat((ang, ang), dim=-1)
            self.cache[key] = (ang.cos().to(dtype), ang.sin().to(dtype))
        return self.cache[key]

    @staticmethod
    def rot(x):
        d = x.shape[-1] // 2
        return torch.cat((-x[..., d:], x[..., :d]), dim=-1)

    def one(self, t, pos, cs, sn):
        c, s = F.embedding(pos, cs)[:, None], F.embedding(pos, sn)[:, None]
        return t * c + self.rot(t) * s

    def forward(self, t, pos):
        d = t.size(-1) // 2
        cs, sn = self.tables(d, int(pos.max()) + 1, t.device, t.dtype)  # the reference's host sync (rope.py:171)
        v, h = t.chunk(2, dim=-1)
        return torch.cat((self.one(v, pos[..., 0], cs, sn), self.one(h, pos[..., 1], cs, sn)), dim=-1)


class Block(nn.Module):
    def __init__(self, qk_norm, rope, eps):
        super().__init__()
        self.norm1, self.norm2 = nn.LayerNorm(C, eps=eps), nn.LayerNorm(C, eps=eps)
        self.qkv, self.proj = nn.Linear(C, 3 * C), nn.Linear(C, C)
        self.q_norm = nn.LayerNorm(HD) if qk_norm else nn.Identity()
        self.k_norm = nn.LayerNorm(HD) if qk_norm else nn.Identity()
        self.fc1, self.fc2, self.act = nn.Linear(C, MLP), nn.Linear(MLP, C), nn.GELU()
        self.g1, self.g2 = nn.Parameter(torch.full((C,), 0.01)), nn.Parameter(torch.full((C,), 0.01))
        self.rope = rope

    def attn(self, x, pos):
        B, N, _ = x.shape
        q, k, v = self.qkv(x).reshape(B, N, 3, HEADS, HD).permute(2, 0, 3, 1, 4).unbind(0)
        q, k = self.q_norm(q), self.k_norm(k)
        if self.rope is not None:
            q, k = self.rope(q, pos), self.rope(k, pos)
        x = F.scaled_dot_product_attention(q, k, v)
        return self.proj(x.transpose(1, 2).reshape(B, N, C))

    def forward(self, x, pos=None):
        x = x + self.attn(self.norm1(x), pos) * self.g1
        return x + self.fc2(self.act(self.fc1(self.norm2(x)))) * self.g2


def model_level(views, iters, depth=24, dino_depth=24, grid=37):
    dev = torch.device("cuda")
    rope = Rope2D()
    with torch.device(dev):
        dino = nn.ModuleList([Block(False, None, 1e-6) for _ in range(dino_depth)])
        frame = nn.ModuleList([Block(True, rope, 1e-5) for _ in range(depth)])
        glob = nn.ModuleList([Block(True, rope, 1e-5) for _ in range(depth)])
    for m in (dino, frame, glob):
        m.eval()
    P, Pd = 7 + grid * grid, 5 + grid * grid
    xd = torch.randn(views, Pd, C, device=dev)
    yy, xx = torch.meshgrid(torch.arange(grid, device=dev), torch.arange(grid, device=dev), indexing="ij")
    pos = torch.cat([torch.zeros(7, 2, dtype=torch.long, device=dev), torch.stack([yy, xx], -1).reshape(-1, 2) + 1], 0)
    pos = pos[None].expand(views, -1, -1).contiguous()

    def dino_fwd():
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            x = xd
            for b in dino:
                x = b(x)
        return x

    def agg_fwd():
        outs = []
        with torch.no_grad(), torch.autocast("cuda", dtype=torch.bfloat16):
            x = torch.cat([xd[:, :5], xd[:, :2], xd[:, 5:]], dim=1)  # [views, P, C]
            for i in range(depth):
                x = frame[i](x.reshape(views, P, C), pos)
                f = x.reshape(1, views, P, C)
                x = glob[i](x.reshape(1, views * P, C), pos.reshape(1, views * P, 2))
                if i in (4, 11, 17, 23):
                    outs.append(torch.cat([f, x.reshape(1, views, P, C)], dim=-1))
        return outs

    t_dino = time_ms(dino_fwd, iters, 2)
    t_agg = time_ms(agg_fwd, iters, 2)
    return {"views": views, "dino_blocks_ms": t_dino, "frame_global_stack_ms": t_agg, "aggregator_ms": t_dino + t_agg,
            "views_per_s": views / ((t_dino + t_agg) * 1e-3),
            "what": "24 DINOv2-shaped blocks + 24 x (frame, global) blocks, torch.autocast(cuda, bf16), random weights; "
                    "patch embedding and prediction heads not included"}


def main():
    ap = argparse.ArgumentParser()
    ap.add_argument("--views", type=int, default=8)
    ap.add_argument("--iters", type=int, default=5)
    ap.add_argument("--out", default="")
    ap.add_argument("--no-model", action="store_true")
    a = ap.parse_args()
    if not torch.cuda.is_available():
        print(json.dumps({"unavailable": "no CUDA device"}))
        return 0
    dev = torch.device("cuda")
    res = {"torch": torch.__version__, "device": torch.cuda.get_device_name(0), "sdpa": [], "linear": []}
    for (b, L, name) in ((1, 11008 * a.views // 8, "global attention"), (a.views, 1376, "frame attention")):
        q, k, v = (torch.randn(b, HEADS, L, HD, device=dev, dtype=torch.bfloat16) for _ in range(3))
        ms = time_ms(lambda: F.scaled_dot_product_attention(q, k, v), 10)
        res["sdpa"].append({"shape": [b, HEADS, L, HD], "what": name, "ms": ms,
                            "tflops": 4.0 * b * L * L * HEADS * HD / ms * 1e-9})
    M = 1376 * a.views
    for (n, k, name) in ((3 * C, C, "qkv"), (C, C, "proj"), (MLP, C, "fc1"), (C, MLP, "fc2")):
        x = torch.randn(M, k, device=dev, dtype=torch.bfloat16)
        w = torch.randn(n, k, device=dev, dtype=torch.bfloat16)
        bias = torch.randn(n, device=dev, dtype=torch.bfloat16)
        ms = time_ms(lambda: F.linear(x, w, bias), 20)
        res["linear"].append({"M": M, "N": n, "K": k, "what": name, "ms": ms, "tflops": 2.0 * M * n * k / ms * 1e-9})
    if not a.no_model:
        res["model"] = model_level(a.views, a.iters)
    line = json.dumps(res)
    print(line)
    if a.out:
        with open(a.out, "w") as f:
            f.write(line + "\n")
    return 0


if __name__ == "__main__":
    sys.exit(main())
