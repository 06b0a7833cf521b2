"""Generate the confidence-filter fixture from the REAL reference (run in the build container only).

    PYTHONPATH=/root/reference:/root/reference/submodules/gsplat python tests/golden/make_golden_conf.py

`GaussianSplatRenderer.apply_confidence_filter` (rasterization.py:248-299) on the seeded clustered splats of
oracle/gs_oracle.py `synthetic_splats` with seeded confidences (1 + exp(.), a few zeros and tiny values that the filter
masks to -inf, all others distinct so that the kept SET is unambiguous): for each (percent, max_gaussians) case the kept
indices (recovered by filtering an index-coded tensor; the reference's torch.topk(sorted=False) order is unspecified, so
they are stored sorted), and for one case the reference's filter -> prune_gs result in full.
"""
import os
import sys

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import gs_oracle as G  # noqa: E402  (only for the seeded inputs)

N_SYN = 12000
CASES = ((30.0, 5000000), (75.0, 5000000), (0.0, 5000000), (99.999, 5000000), (30.0, 1000))
PRUNE_CASE, PRUNE_VOXEL = (30.0, 5000000), 0.01


def synthetic_conf(n: int = N_SYN, seed: int = 9) -> np.ndarray:
    g = np.random.default_rng(seed)
    c = (1.0 + np.exp(g.permutation(np.linspace(-3.0, 3.0, n)))).astype(np.float32)  # distinct in fp32
    c[::173] = 0.0
    c[5::419] = 5e-6
    return c


def main():
    from src.models.models.rasterization import GaussianSplatRenderer  # reference

    conf = synthetic_conf()
    live = conf[conf > 1e-5]
    assert np.unique(live).shape[0] == live.shape[0], "confidences must be distinct"
    syn = G.synthetic_splats(N_SYN)
    tsyn = {k: torch.from_numpy(v)[None] for k, v in syn.items()}
    tsyn["sh"] = tsyn["sh"][:, :, None, :]
    tconf = torch.from_numpy(conf).reshape(1, 2, 60, 100)  # [B, S, H, W]
    coded = {"means": torch.arange(N_SYN, dtype=torch.float32)[None, :, None].expand(1, N_SYN, 3).contiguous()}
    fx = {"conf": conf, "n_syn": N_SYN, "cases": np.array(CASES, dtype=np.float64), "torch": torch.__version__}
    for i, (p, mg) in enumerate(CASES):
        rend = GaussianSplatRenderer(sh_degree=0, predict_offset=False, predict_residual_sh=True, enable_prune=True, voxel_size=0.002,
                                     enable_conf_filter=True, conf_threshold_percent=p, max_gaussians=mg).eval()
        with torch.no_grad():
            idx = rend.apply_confidence_filter(coded, tconf)["means"][0, :, 0].long().numpy()
            fx[f"kept{i}"] = np.sort(idx).astype(np.int32)
            print("percent", p, "max_gaussians", mg, "kept", idx.shape[0])
            if (p, mg) == PRUNE_CASE:
                o = rend.prune_gs(rend.apply_confidence_filter(tsyn, tconf), voxel_size=PRUNE_VOXEL)
                K = o["means"][0].shape[0]
                for k in ("means", "quats", "scales", "opacities", "sh"):
                    fx["pruned_" + k] = o[k][0].reshape(K, -1).numpy()
                print("filter -> prune voxel", PRUNE_VOXEL, "K", K)
    path = os.path.join(HERE, "conf_filter.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()
