"""Generate the rasterisation fixture from gsplat's own pure-torch statement of projection and tile binning (run in the
build container only; gsplat's CUDA extension cannot be built here, its torch functions run on the CPU).

    PYTHONPATH=/root/reference:/root/reference/submodules/gsplat python tests/golden/make_golden_raster.py

Stores, for the seeded scene of oracle/raster_oracle.py `synthetic_scene`: covariances, projected means / conics /
depths / radii (fixed 3.33-sigma extent: the torch statement has no opacity-aware extent), tiles per gaussian, sorted
intersection ids, flatten ids and tile offsets.  Pins oracle/raster_oracle.py `project(opacity_aware=False)` and `bin_tiles`.

Compositing: gsplat's own pure-torch alpha compositing `accumulate` (gsplat/cuda/_torch_impl.py:506-598) is fed with
the (gaussian, pixel, image) index list its CUDA `rasterize_to_indices_in_range` would produce, restated here from
csrc/RasterizeToIndices3DGS.cu:58-160 (per pixel, the tile's gaussians front to back; skip sigma < 0 or alpha < 1/255;
stop, exclusively, when the next transmittance is <= 1e-4).  `accumulate` imports two functions of `nerfacc`, which is
not installed; they are stubbed with their definitions (weights_i = alpha_i * prod_{j<i on the ray} (1 - alpha_j);
accumulate = per-ray sum of weights * values).  Channels: SH-degree-0 colours (rendering.py:509-525) and depth
("RGB+ED").  Stored as acc_renders [C,H,W,4] / acc_alphas [C,H,W,1]: pins oracle/raster_oracle.py `composite`.
"""
import math
import os
import sys
import types

import numpy as np
import torch

HERE = os.path.dirname(os.path.abspath(__file__))
ROOT = os.path.dirname(os.path.dirname(HERE))
sys.path.insert(0, ROOT)
from oracle import raster_oracle as R  # noqa: E402  (only for the seeded inputs)


def install_nerfacc_stub():
    """nerfacc's two functions used by gsplat's `accumulate`, from their definitions (samples of a ray are contiguous)."""
    m = types.ModuleType("nerfacc")

    def render_weight_from_alpha(alphas, ray_indices=None, n_rays=None):
        trans = torch.ones_like(alphas)
        start = torch.ones_like(ray_indices, dtype=torch.bool)
        start[1:] = ray_indices[1:] != ray_indices[:-1]
        run = torch.tensor(1.0, dtype=alphas.dtype)
        for i in range(alphas.shape[0]):  # exclusive product of (1 - alpha) along each ray, in sample order, fp32
            if start[i]:
                run = torch.tensor(1.0, dtype=alphas.dtype)
            trans[i] = run
            run = run * (1.0 - alphas[i])
        return alphas * trans, trans

    def accumulate_along_rays(weights, values=None, ray_indices=None, n_rays=None):
        src = weights[:, None] if values is None else weights[:, None] * values
        out = torch.zeros((n_rays, src.shape[-1]), dtype=src.dtype)
        out.index_add_(0, ray_indices, src)
        return out

    m.render_weight_from_alpha, m.accumulate_along_rays = render_weight_from_alpha, accumulate_along_rays
    sys.modules["nerfacc"] = m


def composite_with_gsplat_accumulate(means2d, conics, depths, opac, sh, flat, offs, W, H):
    from gsplat.cuda._torch_impl import accumulate  # reference dependency

    install_nerfacc_stub()
    Cn, N = depths.shape
    th, tw = offs.shape[-2:]
    m2, con = means2d.reshape(-1, 2), conics.reshape(-1, 3)
    off = offs.reshape(-1).tolist() + [len(flat)]
    g_ids, p_ids, i_ids = [], [], []
    for c in range(Cn):  # csrc/RasterizeToIndices3DGS.cu:58-160, one pixel at a time
        for y in range(H):
            for x in range(W):
                t = (c * th + y // 16) * tw + x // 16
                T = torch.tensor(1.0)
                px, py = torch.tensor(x + 0.5), torch.tensor(y + 0.5)
                for f in flat[off[t]:off[t + 1]].tolist():
                    dx, dy = m2[f, 0] - px, m2[f, 1] - py
                    sigma = 0.5 * (con[f, 0] * dx * dx + con[f, 2] * dy * dy) + con[f, 1] * dx * dy
                    alpha = torch.minimum(torch.tensor(0.999), opac[f % N] * torch.exp(-sigma))
                    if sigma < 0 or alpha < 1.0 / 255.0:
                        continue
                    nT = T * (1.0 - alpha)
                    if nT <= 1e-4:
                        break
                    g_ids.append(f % N), p_ids.append(y * W + x), i_ids.append(c)
                    T = nT
    g_ids, p_ids, i_ids = (torch.tensor(v, dtype=torch.long) for v in (g_ids, p_ids, i_ids))
    colors = torch.clamp_min(0.28209479177387814 * sh + 0.5, 0.0)
    feats = torch.cat([colors[None].expand(Cn, -1, -1), depths[..., None]], dim=-1)  # RGB + depth ("RGB+ED")
    renders, alphas = accumulate(means2d, conics, opac[None].expand(Cn, -1), feats, g_ids, p_ids, i_ids, W, H)
    print("compositing: ", len(g_ids), "(gaussian, pixel) pairs, mean alpha", alphas.mean().item())
    return {"acc_renders": renders.numpy(), "acc_alphas": alphas.numpy(), "acc_pairs": len(g_ids)}


def main():
    from gsplat.cuda._torch_impl import (_fully_fused_projection, _isect_offset_encode, _isect_tiles,  # reference dependency
                                         _quat_scale_to_covar_preci)

    sc = R.synthetic_scene()
    W, H = sc["W"], sc["H"]
    t = {k: torch.from_numpy(v) for k, v in sc.items() if isinstance(v, np.ndarray)}
    viewmats = torch.linalg.inv(t["camtoworlds"])
    covars, _ = _quat_scale_to_covar_preci(t["quats"], t["scales"], True, False, triu=False)
    radii, means2d, depths, conics, _ = _fully_fused_projection(t["means"], covars, viewmats, t["Ks"], W, H, eps2d=0.3,
                                                                near_plane=0.01, far_plane=1e10)
    tw, th = math.ceil(W / 16.0), math.ceil(H / 16.0)
    per, ids, flat = _isect_tiles(means2d, radii, depths, 16, tw, th)
    offs = _isect_offset_encode(ids, viewmats.shape[0], tw, th)
    fx = {"covars": covars.numpy(), "radii": radii.numpy(), "means2d": means2d.numpy(), "depths": depths.numpy(),
          "conics": conics.numpy(), "tiles_per_gauss": per.numpy(), "isect_ids": ids.numpy(), "flatten_ids": flat.numpy(),
          "offsets": offs.numpy(), "viewmats": viewmats.numpy(), "torch": torch.__version__}
    fx.update(composite_with_gsplat_accumulate(means2d, conics, depths, t["opacities"], t["sh"], flat, offs, W, H))
    print("visible (camera, gaussian) pairs", int((radii > 0).all(-1).sum()), "of", radii.shape[0] * radii.shape[1], "intersections", len(ids))
    path = os.path.join(HERE, "raster_scene.npz")
    np.savez_compressed(path, **fx)
    print("wrote", path, os.path.getsize(path) / 1e6, "MB")


if __name__ == "__main__":
    main()
