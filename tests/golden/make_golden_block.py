"""Generate the block-level fixture from the REAL reference modules (run in the build container only).

    PYTHONPATH=/root/reference python tests/golden/make_golden_block.py

SURVEY.md section 8c golden vector (1): the reference `Block` (src/models/layers/block.py) and `Attention`
(src/models/layers/attention.py) in fp32 on the CPU, on seeded tokens [2, 331, 1024] (config C1: 7 special tokens + an
18 x 18 grid) with the aggregator's positions, for the two kinds of block on the path:
  frame  — qk_norm, 2D RoPE (frequency 100), LayerNorm eps 1e-5 (visual_transformer.py:122-141), weights
           `frame_blocks.0.*` of the seeded synthetic state dict;
  dino   — no RoPE, no QK-norm, LayerNorm eps 1e-6 (vision_transformer.py:89), weights `patch_embed.blocks.0.*`.
Both with LayerScale gamma set to 0.5 so that the branches are visible in the output (the synthetic 0.01 hides them).
Stored: every 8th token of each output and the per-token sums of all of them; inputs are regenerated from the seed.
"""
import functools
import importlib
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
W = importlib.import_module("hunyuanworld-mirror_b200.weights")

B, GH, GW, PS, STRIDE, GAMMA = 2, 18, 18, 7, 8, 0.5
KINDS = {"frame": ("frame_blocks.0.", True, True, 1e-5), "dino": ("patch_embed.blocks.0.", False, False, 1e-6)}


def block_inputs(seed: int = 11):
    g = torch.Generator().manual_seed(seed)
    L = PS + GH * GW
    x = torch.randn(B, L, 1024, generator=g) * 0.8
    pos = torch.zeros(B, L, 2, dtype=torch.long)
    i = torch.arange(GH * GW)
    pos[:, PS:, 0], pos[:, PS:, 1] = 1 + i // GW, 1 + i % GW
    return x, pos


def block_weights(pre: str):
    sd = W.synthetic_state_dict(seed=0, depth=1, dino_depth=1)
    sub = {k[len(pre):]: v.clone() for k, v in sd.items() if k.startswith(pre)}
    sub["ls1.gamma"].fill_(GAMMA)
    sub["ls2.gamma"].fill_(GAMMA)
    return sub


def main():
    from src.models.layers.block import Block  # reference
    from src.models.layers.rope import RotaryPositionEmbedding2D  # reference

    x, pos = block_inputs()
    fx = {"stride": STRIDE, "torch": torch.__version__}
    for kind, (pre, qk_norm, rope, eps) in KINDS.items():
        blk = Block(dim=1024, num_heads=16, mlp_ratio=4.0, qkv_bias=True, proj_bias=True, ffn_bias=True, init_values=0.01,
                    norm_layer=functools.partial(torch.nn.LayerNorm, eps=eps), qk_norm=qk_norm,
                    rope=RotaryPositionEmbedding2D(frequency=100) if rope else None).eval()
        blk.load_state_dict(block_weights(pre), strict=True)
        with torch.no_grad():
            out = blk(x, pos=pos if rope else None)
            att = blk.attn(blk.norm1(x), pos=pos if rope else None)
        for name, t in (("block", out), ("attn", att)):
            fx[f"{kind}_{name}_sub"] = t[:, ::STRIDE].numpy()
            fx[f"{kind}_{name}_rowsum"] = t.sum(dim=-1).numpy()
        print(kind, "block |out - x| rel", ((out - x).norm() / x.norm()).item(), "attn abs mean", att.abs().mean().item())
    path = os.path.join(HERE, "block_c1.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()
