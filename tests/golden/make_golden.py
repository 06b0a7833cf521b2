"""Generate golden fixtures from the REAL reference (run in the build container only).

    PYTHONPATH=/root/reference:/root/reference/submodules/gsplat python tests/golden/make_golden.py

Imports `VisualGeometryTransformer` from /root/reference (never copied into this repo), loads
the synthetic state_dict (hunyuanworld-mirror_b200/weights.py) with strict=True, runs the
reference forward in fp32 on CPU (its cuda-autocast regions are no-ops without a GPU) and
stores, per case, a token subset of the four [1,N,P,2048] outputs plus per-token L2 norms of
every token and two stage checkpoints (DINOv2 patch tokens, assembled tokens).  The fixtures
are what pins oracle/aggregator_oracle.py (tests/test_oracle.py) and what the GPU parity
tests compare against at full depth (tests/test_parity_gpu.py).
"""
import importlib
import os
import sys
import time

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
W = importlib.import_module("hunyuanworld-mirror_b200.weights")

CASES = {
    # name: (n_views, H, W, cond_flags, with_priors)
    "c1_nopriors": (2, 252, 252, [0, 0, 0], False),
    "c1_priors": (2, 252, 252, [1, 1, 1], True),
    "rect_partial": (3, 168, 224, [1, 0, 1], True),
    "c2s_nopriors": (2, 518, 518, [0, 0, 0], False),
    # the configuration bench.py times at N = 1 (BASELINE.json configs[1]): global attention over 11 008 tokens
    "c2_nopriors": (8, 518, 518, [0, 0, 0], False),
    # single-GPU stand-in for BASELINE.json configs[2] (pose + depth + intrinsics prior tokens at 518 x 518)
    "c3s_priors": (4, 518, 518, [1, 1, 1], True),
}


def token_subset(P: int, n_views: int = 2) -> np.ndarray:
    """All seven special tokens, the first patches, and patches spread over the grid (fewer per view for many views,
    so that a fixture stays a few MB; the per-token norms stored beside it cover EVERY token)."""
    idx = list(range(0, 9)) + list(np.linspace(9, P - 1, 15 if n_views < 4 else 5).astype(int))
    return np.unique(np.array(idx, dtype=np.int64))


def main():
    from src.models.models.visual_transformer import VisualGeometryTransformer  # reference

    torch.manual_seed(0)
    model = VisualGeometryTransformer(enable_cond=True).eval()
    sd = W.synthetic_state_dict(seed=0)
    missing = model.load_state_dict(sd, strict=True)
    print("loaded", missing)

    stages = {}

    def hook_dino(mod, inp, out):
        stages["dino_patch_tokens"] = out["x_norm_patchtokens"].detach()

    def hook_block0(mod, args, kwargs):
        stages["tokens0"] = args[0].detach()

    model.patch_embed.register_forward_hook(hook_dino)
    model.frame_blocks[0].register_forward_pre_hook(hook_block0, with_kwargs=True)

    only = sys.argv[1:] or list(CASES)
    for name in only:
        n, h, w, flags, with_priors = CASES[name]
        imgs = W.synthetic_images(n, h, w, seed=1)
        priors = W.synthetic_priors(n, h, w, seed=2) if with_priors else None
        t0 = time.time()
        with torch.no_grad():
            outs, psi = model(imgs, priors, cond_flags=flags)
        dt = time.time() - t0
        assert psi == 7 and len(outs) == 4
        P = outs[0].shape[2]
        sel = token_subset(P, n)
        fx = {
            "n_views": n, "H": h, "W": w, "cond_flags": np.array(flags), "with_priors": with_priors,
            "token_idx": sel, "ref_seconds": dt, "torch": torch.__version__,
        }
        for i, o in enumerate(outs):
            assert o.dtype == torch.float32 and tuple(o.shape) == (1, n, P, 2048)
            fx[f"out{i}_sub"] = o[0][:, sel].numpy()
            fx[f"out{i}_norms"] = o[0].norm(dim=-1).numpy()
        g = stages["dino_patch_tokens"].shape[1]
        gsel = np.unique(np.linspace(0, g - 1, 16).astype(int))
        fx["dino_idx"] = gsel
        fx["dino_sub"] = stages["dino_patch_tokens"][:, gsel].numpy()
        fx["dino_norms"] = stages["dino_patch_tokens"].norm(dim=-1).numpy()
        fx["tokens0_sub"] = stages["tokens0"][:, sel].numpy()
        path = os.path.join(HERE, f"{name}.npz")
        np.savez_compressed(path, **fx)
        print(f"{name}: ref forward {dt:.1f}s, P={P}, wrote {os.path.getsize(path) / 1e6:.2f} MB")


if __name__ == "__main__":
    main()
